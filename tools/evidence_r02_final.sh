#!/bin/bash
# Round 2, final evidence run on one B200 (writes into gpurun_out/): GPU test suite, smoke, the bench line, the CPU arm, the ncu
# launch list of the bench command and one full capture of the dominant kernel (each ncu pass after the plain command exited 0).
set -x
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu > gpurun_out/r02_gpu_tests_final.log 2>&1; echo "tests rc=$?"
tail -n 3 gpurun_out/r02_gpu_tests_final.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02_smoke_final.log 2>&1; echo "smoke rc=$?"
python bench.py > gpurun_out/r02_bench_v5.json 2> gpurun_out/r02_bench_v5.err; echo "bench rc=$?"
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r02_bench_reference_v3.json 2> gpurun_out/r02_bench_reference_v3.err; echo "reference rc=$?"
python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/r02_bench_short.json 2>/dev/null && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/r02_launches_bench_v5.csv python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/ncu_launches.log 2>&1; echo "launch list rc=$?"
python tools/bench_pipeline.py --bytes 33554432 --no-cpu > gpurun_out/r02_pipeline_32MiB.json 2>/dev/null && \
ncu --set full --clock-control none --import-source on -k regex:"k_eval_count_join|k_eval_write" -c 4 -o gpurun_out/eval_v11 python tools/bench_pipeline.py --bytes 33554432 --no-cpu > gpurun_out/ncu_v11.log 2>&1; echo "ncu full rc=$?"
tail -c 400 gpurun_out/r02_bench_v5.json
