#!/bin/bash
# Round 2, thirtieth GPU call: which child the exporting thread keeps — reflection, refraction, or the one that stays inside the medium.
cd "$(dirname "$0")/.."
O=gpurun_out
export VARIANT_SHARE=8 VARIANT_STEPS=15 VARIANT_TIMEOUT=240
VARIANT_SCENES=sponza,dielectrics,dynamic timeout 1200 python tools/variant_bench.py "off,queue=off" "all_keep0,queue=all,keep=0" "all_keep1,queue=all,keep=1" "all_keep2,queue=all,keep=2" "glass_keep2,queue=glass,keep=2" > $O/c30_variants.jsonl 2> $O/c30_variants.err
python - <<'PY'
import json
for l in open("gpurun_out/c30_variants.jsonl"):
    d = json.loads(l)
    print(d.get("spec"), d.get("error", "")[:200], d.get("parity_failures"), {k: (d[k]["ms_render"], d[k]["tail_ms"], d[k]["ms_render_share8"], d[k]["tail_ms_share8"]) for k in ("sponza", "dielectrics", "dynamic") if k in d})
PY
B200RT_QUEUE_KEEP_REFRACTION=2 STRESS_TIMEOUT=200 timeout 900 python tools/queue_stress.py 500 > $O/c30_queue_stress_keep2.jsonl 2> $O/c30_queue_stress.err; cat $O/c30_queue_stress_keep2.jsonl
