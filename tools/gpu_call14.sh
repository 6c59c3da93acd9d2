#!/bin/bash
# Round 2, fourteenth GPU call: everything at the head of the round — suite, smoke, stress, bench lines, then the profiler passes.
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
rm -f $O/full_size_reference.jsonl
timeout 900 python -m pytest tests -m gpu -x -q > $O/c14_gpu_tests.log 2>&1; echo "pytest rc $?" >> $O/c14_gpu_tests.log
tail -3 $O/c14_gpu_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/c14_smoke.log 2>&1; echo "smoke rc $?" >> $O/c14_smoke.log; tail -2 $O/c14_smoke.log
STRESS_TIMEOUT=200 timeout 1300 python tools/queue_stress.py 1000 > $O/c14_queue_stress.jsonl 2> $O/c14_queue_stress.err; cat $O/c14_queue_stress.jsonl
timeout 400 python bench.py --steps 20 --warmup 3 > $O/c14_bench.json 2> $O/c14_bench.err; echo "bench rc $?" >> $O/c14_bench.err
timeout 400 python bench.py --steps 10 --warmup 3 --workload dielectrics > $O/c14_bench_dielectrics.json 2> $O/c14_bench_dielectrics.err; echo "bench rc $?" >> $O/c14_bench_dielectrics.err
timeout 400 python bench.py --steps 10 --warmup 3 --workload sibenik4k > $O/c14_bench_sibenik4k.json 2> $O/c14_bench_sibenik4k.err; echo "bench rc $?" >> $O/c14_bench_sibenik4k.err
timeout 400 python bench.py --steps 20 --warmup 3 --arithmetic fast > $O/c14_bench_fast.json 2> $O/c14_bench_fast.err; echo "bench rc $?" >> $O/c14_bench_fast.err
timeout 300 python bench.py --impl reference --steps 5 --warmup 1 > $O/c14_bench_reference.json 2> $O/c14_bench_reference.err; echo "bench rc $?" >> $O/c14_bench_reference.err
tail -n 1 $O/c14_bench*.err; cut -c1-260 $O/c14_bench.json
if grep -q "bench rc 0" $O/c14_bench.err; then
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/c14_launches_bench.csv python bench.py --steps 2 --warmup 3 > $O/c14_ncu_launches.log 2>&1
  EXTRA=lts__t_bytes.sum,l1tex__t_bytes.sum,smsp__thread_inst_executed.sum
  for W in sponza1080 dielectrics sibenik4k; do
    B200RT_DEEP_QUEUE=off timeout 600 ncu --set full --metrics $EXTRA --clock-control none --import-source on -k regex:render_kernel -s 3 -c 1 -f -o $O/c14_render_${W} python tools/ncu_frame.py $W > $O/c14_ncu_${W}.log 2>&1
  done
  B200RT_DEEP_QUEUE=all timeout 600 ncu --set full --metrics $EXTRA --clock-control none --import-source on -k regex:render_queue_kernel -s 3 -c 1 -f -o $O/c14_render_queue_dielectrics python tools/ncu_frame.py dielectrics > $O/c14_ncu_queue_dielectrics.log 2>&1
  B200RT_DEEP_QUEUE=off timeout 600 ncu --set full --metrics $EXTRA --clock-control none --import-source on -k regex:fxaa_kernel -s 3 -c 1 -f -o $O/c14_fxaa_sponza1080 python tools/ncu_frame.py sponza1080 > $O/c14_ncu_fxaa.log 2>&1
  ls -la $O/c14*.ncu-rep
fi
