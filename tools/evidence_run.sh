#!/bin/bash
# One evidence pass on the GPU box: tests, bench lines, launch list and ncu captures (each capture only after the same
# command has exited 0 without ncu). Results under gpurun_out/; the summaries that are judged are copied to profiles/.
set -u
TAG=${1:-v10}
O=gpurun_out
mkdir -p $O
python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/${TAG}_pytest_gpu.log
python bench.py --impl reference --steps 2 --warmup 1 > $O/${TAG}_bench_reference.json 2> $O/${TAG}_bench_reference.err; echo "reference rc=$?"
python bench.py --steps 5 --warmup 3 > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err; echo "bench rc=$?"
python tools/bench_matcher.py --docs 1000 --tokens 1000 --rules 1000 > $O/${TAG}_matcher.json 2> $O/${TAG}_matcher.err; echo "matcher rc=$?"
python tools/bench_matcher.py --docs 8000 --tokens 1000 --rules 1000 --no-ref > $O/${TAG}_matcher_8000docs.json 2>> $O/${TAG}_matcher.err; echo "matcher8000 rc=$?"
python tools/bench_pipeline.py > $O/${TAG}_pipeline_c3.json 2> $O/${TAG}_pipeline_c3.err; echo "pipeline rc=$?"
python bench.py --steps 1 --warmup 1 --no-cpu --bytes 268435456 > $O/${TAG}_plain256.json 2>&1 && {
  ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${TAG}_launches_256MiB.csv python bench.py --steps 1 --warmup 1 --no-cpu --bytes 268435456 > $O/${TAG}_ncu_launches.log 2>&1
  ncu --set full --clock-control none --import-source on -k regex:"k_lex_resolve_chunks|k_lex_emit|k_lex_count|k_lex_finish|k_fused_write" -c 5 -o $O/r01_lexer_${TAG} -f python bench.py --steps 1 --warmup 1 --no-cpu --bytes 268435456 > $O/${TAG}_ncu_lexer.log 2>&1
}
python tools/bench_matcher.py --docs 1000 --tokens 1000 --rules 1000 --steps 1 --no-ref > $O/${TAG}_matcher_plain.json 2>&1 && \
  ncu --set full --clock-control none --import-source on -k regex:k_match_documents -c 1 -o $O/r01_matcher_${TAG} -f python tools/bench_matcher.py --docs 1000 --tokens 1000 --rules 1000 --steps 1 --no-ref > $O/${TAG}_ncu_matcher.log 2>&1
echo done; tail -1 $O/${TAG}_bench.json | cut -c1-300
