#!/bin/bash
# Round 2, seventh GPU call: the no-alloc entry point's test alone, the queue under stress, the whole gpu suite, smoke, bench lines,
# and the queued kernel re-captured (one queue now).
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
rm -f $O/full_size_reference.jsonl
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > $O/c7_no_alloc_alone.log 2>&1; echo "pytest rc $?" >> $O/c7_no_alloc_alone.log
tail -3 $O/c7_no_alloc_alone.log
STRESS_TIMEOUT=200 timeout 1300 python tools/queue_stress.py 1000 > $O/c7_queue_stress.jsonl 2> $O/c7_queue_stress.err
cat $O/c7_queue_stress.jsonl
timeout 900 python -m pytest tests -m gpu -x -q > $O/c7_gpu_tests.log 2>&1; echo "pytest rc $?" >> $O/c7_gpu_tests.log
tail -4 $O/c7_gpu_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/c7_smoke.log 2>&1; echo "smoke rc $?" >> $O/c7_smoke.log; tail -2 $O/c7_smoke.log
timeout 400 python bench.py --steps 20 --warmup 3 > $O/c7_bench.json 2> $O/c7_bench.err; echo "bench rc $?" >> $O/c7_bench.err
timeout 400 python bench.py --steps 10 --warmup 3 --workload dielectrics > $O/c7_bench_dielectrics.json 2> $O/c7_bench_dielectrics.err; echo "bench rc $?" >> $O/c7_bench_dielectrics.err
tail -n 2 $O/c7_bench.err $O/c7_bench_dielectrics.err; cut -c1-300 $O/c7_bench.json $O/c7_bench_dielectrics.json
export VARIANT_SHARE=8 VARIANT_STEPS=10 VARIANT_TIMEOUT=150
timeout 600 python tools/variant_bench.py "off,queue=off" "auto,queue=auto" "queue_all,queue=all" "queue_glass,queue=glass" > $O/c7_variants.jsonl 2> $O/c7_variants.err
VARIANT_SCENES=dynamic timeout 300 python tools/variant_bench.py "off,queue=off" "queue_all,queue=all" "queue_glass,queue=glass" > $O/c7_variants_dynamic.jsonl 2>> $O/c7_variants.err
cut -c1-600 $O/c7_variants.jsonl $O/c7_variants_dynamic.jsonl
EXTRA=lts__t_bytes.sum,l1tex__t_bytes.sum,smsp__thread_inst_executed.sum
B200RT_DEEP_QUEUE=all timeout 600 ncu --set full --metrics $EXTRA --clock-control none --import-source on -k regex:render_queue_kernel -s 3 -c 1 -f -o $O/c7_render_queue_dielectrics python tools/ncu_frame.py dielectrics > $O/c7_ncu_queue_dielectrics.log 2>&1
ls -la $O/*.ncu-rep | tail -2
