#!/bin/bash
# Round 2, fifth GPU call: the whole gpu suite (full-size rows incl. stock / queue / FAST, the reference through the bridge),
# (Historical: as run. The defer= / B200RT_QUEUE_DEFER_SHADOWS settings it names belonged to the shadow-ray queue, which was removed afterwards.)
# queue settings side by side, the bench lines, then — only after those exited — the ncu launch list and full captures.
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
rm -f $O/full_size_reference.jsonl
nvidia-smi --query-gpu=name,clocks.max.sm,memory.total --format=csv > $O/c5_box.txt 2>&1; nproc >> $O/c5_box.txt

timeout 1500 python -m pytest tests -m gpu -x -q > $O/c5_gpu_tests.log 2>&1; echo "pytest rc $?" >> $O/c5_gpu_tests.log
tail -4 $O/c5_gpu_tests.log

export VARIANT_SHARE=8 VARIANT_STEPS=10 VARIANT_TIMEOUT=150
timeout 900 python tools/variant_bench.py "off,queue=off" "auto,queue=auto" "queue_all_shadows_inline,queue=all,defer=0" "queue_glass_shadows_inline,queue=glass,defer=0" "queue_all_deferred,queue=all,defer=1" \
  "fast_off,queue=off,arith=fast" "fast_auto,queue=auto,arith=fast" > $O/c5_variants.jsonl 2> $O/c5_variants.err
VARIANT_SCENES=sibenik,dynamic timeout 600 python tools/variant_bench.py "off,queue=off" "auto,queue=auto" "fast_auto,queue=auto,arith=fast" > $O/c5_variants_other.jsonl 2> $O/c5_variants_other.err
cut -c1-420 $O/c5_variants.jsonl $O/c5_variants_other.jsonl

timeout 400 python bench.py --steps 20 --warmup 3 > $O/c5_bench.json 2> $O/c5_bench.err; echo "bench rc $?" >> $O/c5_bench.err
timeout 400 python bench.py --steps 10 --warmup 3 --workload dielectrics > $O/c5_bench_dielectrics.json 2> $O/c5_bench_dielectrics.err; echo "bench rc $?" >> $O/c5_bench_dielectrics.err
timeout 400 python bench.py --steps 10 --warmup 3 --workload sibenik4k > $O/c5_bench_sibenik4k.json 2> $O/c5_bench_sibenik4k.err; echo "bench rc $?" >> $O/c5_bench_sibenik4k.err
timeout 400 python bench.py --steps 20 --warmup 3 --arithmetic fast > $O/c5_bench_fast.json 2> $O/c5_bench_fast.err; echo "bench rc $?" >> $O/c5_bench_fast.err
timeout 300 python bench.py --impl reference --steps 5 --warmup 1 > $O/c5_bench_reference.json 2> $O/c5_bench_reference.err; echo "bench rc $?" >> $O/c5_bench_reference.err
tail -2 $O/c5_bench*.err; cut -c1-600 $O/c5_bench.json

# ---- profiler passes (numbers printed under ncu are never bench values) ----
if grep -q "bench rc 0" $O/c5_bench.err; then
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/c5_launches_bench.csv python bench.py --steps 2 --warmup 3 --no-scale-4k > $O/c5_ncu_launches.log 2>&1
  EXTRA=lts__t_bytes.sum,l1tex__t_bytes.sum,smsp__thread_inst_executed.sum
  for W in sponza1080 dielectrics sibenik4k; do
    B200RT_DEEP_QUEUE=off timeout 600 ncu --set full --metrics $EXTRA --clock-control none --import-source on -k regex:render_kernel -s 3 -c 1 -f -o $O/c5_render_${W} python tools/ncu_frame.py $W > $O/c5_ncu_${W}.log 2>&1
  done
  B200RT_DEEP_QUEUE=all B200RT_QUEUE_DEFER_SHADOWS=0 timeout 600 ncu --set full --metrics $EXTRA --clock-control none --import-source on -k regex:render_queue_kernel -s 3 -c 1 -f -o $O/c5_render_queue_dielectrics python tools/ncu_frame.py dielectrics > $O/c5_ncu_queue_dielectrics.log 2>&1
  B200RT_DEEP_QUEUE=off timeout 600 ncu --set full --metrics $EXTRA --clock-control none --import-source on -k regex:fxaa_kernel -s 3 -c 1 -f -o $O/c5_fxaa_sponza1080 python tools/ncu_frame.py sponza1080 > $O/c5_ncu_fxaa.log 2>&1
  ls -la $O/*.ncu-rep
fi
