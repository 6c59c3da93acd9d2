#!/bin/bash
# Round 2, twenty-fifth GPU call: how many items must wait before a warp prefers them to a tile unit.
cd "$(dirname "$0")/.."
O=gpurun_out
export VARIANT_SHARE=8 VARIANT_STEPS=15 VARIANT_TIMEOUT=240
VARIANT_SCENES=sponza,dielectrics,dynamic timeout 1200 python tools/variant_bench.py "all_batch32,queue=all" "all_batch1,queue=all,batch=1" "all_batch256,queue=all,batch=256" "all_batch2048,queue=all,batch=2048" "all_batch1000000,queue=all,batch=1000000" > $O/c25_variants.jsonl 2> $O/c25_variants.err
python - <<'PY'
import json
for l in open("gpurun_out/c25_variants.jsonl"):
    d = json.loads(l)
    print(d.get("spec"), d.get("error", "")[:200], d.get("parity_failures"), {k: (d[k]["ms_render"], d[k]["tail_ms"], d[k]["ms_render_share8"], d[k]["tail_ms_share8"]) for k in ("sponza", "dielectrics", "dynamic") if k in d})
PY
