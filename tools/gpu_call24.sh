#!/bin/bash
# Round 2, twenty-fourth GPU call: after grouping the queue slots by ray kind — suite, stress, bench lines.
cd "$(dirname "$0")/.."
O=gpurun_out
rm -f $O/full_size_reference.jsonl
timeout 900 python -m pytest tests -m gpu -x -q > $O/c24_gpu_tests.log 2>&1; echo "pytest rc $?" >> $O/c24_gpu_tests.log
tail -3 $O/c24_gpu_tests.log
STRESS_TIMEOUT=200 timeout 1300 python tools/queue_stress.py 1000 > $O/c24_queue_stress.jsonl 2> $O/c24_queue_stress.err; cat $O/c24_queue_stress.jsonl
timeout 400 python bench.py --steps 20 --warmup 3 > $O/c24_bench.json 2> $O/c24_bench.err; echo "bench rc $?" >> $O/c24_bench.err
timeout 400 python bench.py --steps 10 --warmup 3 --workload dielectrics > $O/c24_bench_dielectrics.json 2> $O/c24_bench_dielectrics.err; echo "bench rc $?" >> $O/c24_bench_dielectrics.err
tail -n 1 $O/c24_bench.err $O/c24_bench_dielectrics.err; cut -c1-250 $O/c24_bench.json $O/c24_bench_dielectrics.json
EXTRA=lts__t_bytes.sum,l1tex__t_bytes.sum,smsp__thread_inst_executed.sum
B200RT_DEEP_QUEUE=all timeout 600 ncu --set full --metrics $EXTRA --clock-control none --import-source on -k regex:render_queue_kernel -s 5 -c 1 -f -o $O/c24_render_queue_dielectrics python tools/ncu_frame.py dielectrics > $O/c24_ncu_queue_dielectrics.log 2>&1
ls -la $O/c24*.ncu-rep
