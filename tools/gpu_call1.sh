#!/bin/bash
# Round 2, first GPU call: sanity of the restructured library, then every prepared variant / setting timed side by side.
# Runs ON the GPU box (gpurun -- bash tools/gpu_call1.sh); everything lands in gpurun_out/.
cd "$(dirname "$0")/.."
O=gpurun_out
B=cpu-raytracer_b200/build
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > $O/c1_gpu.txt 2>&1
# 1. parity of the shipped library: the fast part of the GPU suite (exact mode, queue off = default)
timeout 300 python -m pytest tests/test_device_parity.py tests/test_device_edges.py tests/test_config_switches.py tests/test_texture_ingest.py -m gpu -x -q > $O/c1_tests.log 2>&1; echo "tests rc $?" >> $O/c1_tests.log
# 2. one small queued frame under the memory checker
timeout 240 compute-sanitizer --tool memcheck --error-exitcode 9 python tools/queue_smoke.py > $O/c1_memcheck.log 2>&1; echo "memcheck rc $?" >> $O/c1_memcheck.log
# 3. variants and settings
export VARIANT_SHARE=8 VARIANT_STEPS=10 VARIANT_TIMEOUT=200
timeout 1500 python tools/variant_bench.py \
  "shipped" \
  "queue_glass,queue=glass" "queue_glass_inline-1,queue=glass,inline=-1" "queue_glass_inline1,queue=glass,inline=1" "queue_all,queue=all" \
  "octant,lib=$B/libb200rt_octant.so" "prefetch,lib=$B/libb200rt_prefetch.so" "octant_prefetch,lib=$B/libb200rt_octant_prefetch.so" "wide4,lib=$B/libb200rt_wide4.so" \
  "fast,arith=fast" "fast_contraction_only,lib=$B/libb200rt_fast0.so,arith=fast" "fast_slab_only,lib=$B/libb200rt_fast1.so,arith=fast" \
  "fast_triangle_only,lib=$B/libb200rt_fast2.so,arith=fast" "fast_texel_only,lib=$B/libb200rt_fast4.so,arith=fast" \
  "fast_queue_glass,arith=fast,queue=glass" \
  > $O/c1_variants.jsonl 2> $O/c1_variants.err
# 4. the other two scenes for the settings that matter
VARIANT_SCENES=sibenik,dynamic timeout 600 python tools/variant_bench.py "shipped" "queue_glass,queue=glass" "queue_all,queue=all" "fast,arith=fast" > $O/c1_variants_other.jsonl 2> $O/c1_variants_other.err
# 5. the bench line as it stands
timeout 300 python bench.py --steps 10 --warmup 3 > $O/c1_bench.json 2> $O/c1_bench.err
tail -3 $O/c1_tests.log $O/c1_memcheck.log; cat $O/c1_variants.jsonl $O/c1_variants_other.jsonl | cut -c1-900
