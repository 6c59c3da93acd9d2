#!/bin/bash
# Round 2, third GPU call: the deep-pixel queue with ticket claims and the export hook; the new bench line; the 2-process peer exchange.
cd "$(dirname "$0")/.."
O=gpurun_out
export VARIANT_SHARE=8 VARIANT_STEPS=10 VARIANT_TIMEOUT=120
timeout 700 python tools/variant_bench.py "shipped" \
  "queue_glass,queue=glass" "queue_glass_inline1,queue=glass,inline=1" "queue_glass_inline2,queue=glass,inline=2" "queue_all,queue=all" \
  > $O/c3_variants.jsonl 2> $O/c3_variants.err
B200RT_QUEUE_DEFER_SHADOWS=0 timeout 300 python tools/variant_bench.py "queue_glass_shadows_inline,queue=glass" "queue_all_shadows_inline,queue=all" >> $O/c3_variants.jsonl 2>> $O/c3_variants.err
VARIANT_SCENES=sibenik,dynamic timeout 500 python tools/variant_bench.py "shipped" "queue_glass,queue=glass" "queue_all,queue=all" > $O/c3_variants_other.jsonl 2> $O/c3_variants_other.err
timeout 200 python -m pytest tests/test_peer_exchange.py -m gpu -x -q > $O/c3_peer.log 2>&1; echo "peer rc $?" >> $O/c3_peer.log
timeout 400 python bench.py --steps 5 --warmup 3 > $O/c3_bench.json 2> $O/c3_bench.err; echo "bench rc $?" >> $O/c3_bench.err
tail -5 $O/c3_peer.log; tail -3 $O/c3_bench.err; cat $O/c3_variants.jsonl $O/c3_variants_other.jsonl | cut -c1-300
