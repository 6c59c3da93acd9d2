#!/bin/bash
# Round 2, twenty-eighth GPU call: bench lines and the queued kernel's capture at the head of the round.
cd "$(dirname "$0")/.."
O=gpurun_out
timeout 400 python bench.py --steps 20 --warmup 3 > $O/c28_bench.json 2> $O/c28_bench.err; echo "bench rc $?" >> $O/c28_bench.err
timeout 400 python bench.py --steps 10 --warmup 3 --workload dielectrics > $O/c28_bench_dielectrics.json 2> $O/c28_bench_dielectrics.err; echo "bench rc $?" >> $O/c28_bench_dielectrics.err
timeout 400 python bench.py --steps 10 --warmup 3 --workload sibenik4k > $O/c28_bench_sibenik4k.json 2> $O/c28_bench_sibenik4k.err; echo "bench rc $?" >> $O/c28_bench_sibenik4k.err
timeout 400 python bench.py --steps 20 --warmup 3 --arithmetic fast > $O/c28_bench_fast.json 2> $O/c28_bench_fast.err; echo "bench rc $?" >> $O/c28_bench_fast.err
timeout 400 python bench.py --steps 10 --warmup 3 --arithmetic fast --workload dielectrics > $O/c28_bench_fast_dielectrics.json 2> $O/c28_bench_fast_dielectrics.err; echo "bench rc $?" >> $O/c28_bench_fast_dielectrics.err
tail -n 1 $O/c28_bench*.err; cut -c1-250 $O/c28_bench.json $O/c28_bench_dielectrics.json
EXTRA=lts__t_bytes.sum,l1tex__t_bytes.sum,smsp__thread_inst_executed.sum
B200RT_DEEP_QUEUE=all timeout 600 ncu --set full --metrics $EXTRA --clock-control none --import-source on -k regex:render_queue_kernel -s 5 -c 1 -f -o $O/c28_render_queue_dielectrics python tools/ncu_frame.py dielectrics > $O/c28_ncu_queue_dielectrics.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/c28_launches_bench.csv python bench.py --steps 2 --warmup 3 > $O/c28_ncu_launches.log 2>&1
ls -la $O/c28*.ncu-rep $O/c28_launches_bench.csv
