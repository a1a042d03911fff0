#!/bin/bash
# Round 2 evidence runs on the GPU box (writes into gpurun_out/): level-2 timings on the C1 and C4 shapes with parity
# checks, small batch latencies, compute-sanitizer memcheck / racecheck on the smoke run and on evaluator + lexer tests.
set -x
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 300 python tools/bench_matcher.py --docs 1000 --check > gpurun_out/r02_matcher_c1_1000docs.json 2> gpurun_out/ev.err
timeout 300 python tools/bench_matcher.py --docs 8000 > gpurun_out/r02_matcher_c1_8000docs.json 2>> gpurun_out/ev.err
timeout 600 python tools/bench_matcher.py --c4 --rules 100000 --features 50000 --docs 4800 > gpurun_out/r02_matcher_c4_4800docs.json 2>> gpurun_out/ev.err
timeout 600 python tools/bench_matcher.py --c4 --rules 100000 --features 50000 --docs 400 --check > gpurun_out/r02_matcher_c4_400docs.json 2>> gpurun_out/ev.err
timeout 300 python tools/bench_small_batches.py > gpurun_out/r02_small_batches.json 2>> gpurun_out/ev.err
for tool in memcheck racecheck; do
  timeout 900 compute-sanitizer --tool $tool --print-limit 10 --log-file gpurun_out/r02_sanitizer_${tool}_smoke.txt python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/san_${tool}_smoke.out 2>&1
  timeout 900 compute-sanitizer --tool $tool --print-limit 10 --log-file gpurun_out/r02_sanitizer_${tool}_tests.txt python -m pytest -x -q "tests/test_matcher_eval_gpu.py::test_nested_any_expressions" "tests/test_matcher_gpu.py::test_c1_style_rules" "tests/test_lexer_gpu.py::test_golden_case1_exact" > gpurun_out/san_${tool}_tests.out 2>&1
done
tail -2 gpurun_out/san_*_tests.out gpurun_out/san_*_smoke.out
tail -4 gpurun_out/r02_sanitizer_*.txt
cat gpurun_out/r02_matcher_*.json gpurun_out/r02_small_batches.json
