#!/bin/bash
# Round 2, ninth GPU call: after "one queue slot per child ray" and the smaller TLAS stack — suite, queue under stress, variants, bench lines.
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
rm -f $O/full_size_reference.jsonl
timeout 900 python -m pytest tests -m gpu -x -q > $O/c9_gpu_tests.log 2>&1; echo "pytest rc $?" >> $O/c9_gpu_tests.log
tail -3 $O/c9_gpu_tests.log
STRESS_TIMEOUT=200 timeout 1300 python tools/queue_stress.py 1000 > $O/c9_queue_stress.jsonl 2> $O/c9_queue_stress.err
cat $O/c9_queue_stress.jsonl
export VARIANT_SHARE=8 VARIANT_STEPS=10 VARIANT_TIMEOUT=150
timeout 600 python tools/variant_bench.py "off,queue=off" "auto,queue=auto" "queue_all,queue=all" "queue_glass,queue=glass" "fast_auto,queue=auto,arith=fast" > $O/c9_variants.jsonl 2> $O/c9_variants.err
VARIANT_SCENES=sibenik,dynamic timeout 400 python tools/variant_bench.py "off,queue=off" "auto,queue=auto" "queue_all,queue=all" > $O/c9_variants_other.jsonl 2>> $O/c9_variants.err
cut -c1-600 $O/c9_variants.jsonl $O/c9_variants_other.jsonl
timeout 400 python bench.py --steps 20 --warmup 3 > $O/c9_bench.json 2> $O/c9_bench.err; echo "bench rc $?" >> $O/c9_bench.err
timeout 400 python bench.py --steps 10 --warmup 3 --workload dielectrics > $O/c9_bench_dielectrics.json 2> $O/c9_bench_dielectrics.err; echo "bench rc $?" >> $O/c9_bench_dielectrics.err
tail -n 1 $O/c9_bench.err $O/c9_bench_dielectrics.err; cut -c1-250 $O/c9_bench.json $O/c9_bench_dielectrics.json
