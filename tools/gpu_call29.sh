#!/bin/bash
# Round 2, last GPU call: the clean rebuild — suite, smoke, the bench line.
cd "$(dirname "$0")/.."
O=gpurun_out
rm -f $O/full_size_reference.jsonl
timeout 900 python -m pytest tests -m gpu -x -q > $O/c29_gpu_tests.log 2>&1; echo "pytest rc $?" >> $O/c29_gpu_tests.log
tail -3 $O/c29_gpu_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/c29_smoke.log 2>&1; echo "smoke rc $?" >> $O/c29_smoke.log; tail -2 $O/c29_smoke.log | cut -c1-300
timeout 400 python bench.py > $O/c29_bench.json 2> $O/c29_bench.err; echo "bench rc $?" >> $O/c29_bench.err; tail -1 $O/c29_bench.err; cut -c1-300 $O/c29_bench.json
timeout 300 python bench.py --impl reference --steps 5 --warmup 1 > $O/c29_bench_reference.json 2> $O/c29_bench_reference.err; echo "bench rc $?" >> $O/c29_bench_reference.err; tail -1 $O/c29_bench_reference.err; cut -c1-200 $O/c29_bench_reference.json
