#!/bin/bash
# Round 2, thirty-first GPU call: "the child inside the medium stays" as the shipped setting — suite, stress, bench lines, the queued kernel's capture.
cd "$(dirname "$0")/.."
O=gpurun_out
rm -f $O/full_size_reference.jsonl
timeout 900 python -m pytest tests -m gpu -x -q > $O/c31_gpu_tests.log 2>&1; echo "pytest rc $?" >> $O/c31_gpu_tests.log
tail -3 $O/c31_gpu_tests.log
STRESS_TIMEOUT=200 timeout 1300 python tools/queue_stress.py 1000 > $O/c31_queue_stress.jsonl 2> $O/c31_queue_stress.err; cat $O/c31_queue_stress.jsonl
timeout 400 python bench.py --steps 20 --warmup 3 > $O/c31_bench.json 2> $O/c31_bench.err; echo "bench rc $?" >> $O/c31_bench.err
timeout 400 python bench.py --steps 10 --warmup 3 --workload dielectrics > $O/c31_bench_dielectrics.json 2> $O/c31_bench_dielectrics.err; echo "bench rc $?" >> $O/c31_bench_dielectrics.err
timeout 400 python bench.py --steps 10 --warmup 3 --arithmetic fast --workload dielectrics > $O/c31_bench_fast_dielectrics.json 2> $O/c31_bench_fast_dielectrics.err; echo "bench rc $?" >> $O/c31_bench_fast_dielectrics.err
tail -n 1 $O/c31_bench*.err; cut -c1-250 $O/c31_bench.json $O/c31_bench_dielectrics.json
EXTRA=lts__t_bytes.sum,l1tex__t_bytes.sum,smsp__thread_inst_executed.sum
B200RT_DEEP_QUEUE=all timeout 600 ncu --set full --metrics $EXTRA --clock-control none --import-source on -k regex:render_queue_kernel -s 5 -c 1 -f -o $O/c31_render_queue_dielectrics python tools/ncu_frame.py dielectrics > $O/c31_ncu_queue_dielectrics.log 2>&1
ls -la $O/c31*.ncu-rep
