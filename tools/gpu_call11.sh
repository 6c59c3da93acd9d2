#!/bin/bash
# Round 2, eleventh GPU call: how long idle warps keep looking at the queue, and how many levels a popped ray walks itself.
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
export VARIANT_SHARE=8 VARIANT_STEPS=15 VARIANT_TIMEOUT=200
timeout 1500 python tools/variant_bench.py "off,queue=off" "all_looks3,queue=all,looks=3" "all_looks10,queue=all,looks=10" "all_looks100,queue=all,looks=100" "all_looks1000,queue=all,looks=1000" \
   "all_inline1,queue=all,inline=1" "all_inline2,queue=all,inline=2" "all_inline1_looks100,queue=all,inline=1,looks=100" "glass_looks100,queue=glass,looks=100" "glass_inline1_looks100,queue=glass,inline=1,looks=100" > $O/c11_variants.jsonl 2> $O/c11_variants.err
python - <<'PY'
import json
for l in open("gpurun_out/c11_variants.jsonl"):
    d = json.loads(l)
    print(d.get("spec"), d.get("error", "")[:200], d.get("parity_failures"), {k: (d[k]["ms_render"], d[k]["tail_ms"], d[k]["ms_render_share8"], d[k]["tail_ms_share8"]) for k in ("sponza", "dielectrics") if k in d})
PY
