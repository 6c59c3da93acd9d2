#!/bin/bash
# Round 2, second GPU call: the deep-pixel queue after the polling fix, all settings, four scenes.
cd "$(dirname "$0")/.."
O=gpurun_out
timeout 300 python -m pytest tests/test_device_parity.py tests/test_device_edges.py -m gpu -x -q > $O/c2_tests.log 2>&1; echo "tests rc $?" >> $O/c2_tests.log
export VARIANT_SHARE=8 VARIANT_STEPS=10 VARIANT_TIMEOUT=200
timeout 900 python tools/variant_bench.py "shipped" \
  "queue_glass,queue=glass" "queue_glass_inline-1,queue=glass,inline=-1" "queue_glass_inline1,queue=glass,inline=1" "queue_glass_inline2,queue=glass,inline=2" \
  "queue_all,queue=all" "queue_all_inline1,queue=all,inline=1" "fast_queue_glass,arith=fast,queue=glass" \
  > $O/c2_variants.jsonl 2> $O/c2_variants.err
VARIANT_SCENES=sibenik,dynamic timeout 600 python tools/variant_bench.py "shipped" "queue_glass,queue=glass" "queue_glass_inline1,queue=glass,inline=1" "queue_all,queue=all" > $O/c2_variants_other.jsonl 2> $O/c2_variants_other.err
tail -3 $O/c2_tests.log; cat $O/c2_variants.jsonl $O/c2_variants_other.jsonl | cut -c1-1200
