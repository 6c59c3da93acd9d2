#!/bin/bash
# Round 2, eighteenth GPU call: idle warps that stay longer but look rarely (pause doubling up to 32 / 128 us).
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
export VARIANT_SHARE=8 VARIANT_STEPS=15 VARIANT_TIMEOUT=240
timeout 1200 python tools/variant_bench.py "off,queue=off" "all,queue=all" "all_looks20_shift4,queue=all,looks=20,shift=4" "all_looks40_shift4,queue=all,looks=40,shift=4" "all_looks20_shift6,queue=all,looks=20,shift=6" "all_looks100_shift5,queue=all,looks=100,shift=5" \
   "glass_looks40_shift4,queue=glass,looks=40,shift=4" > $O/c18_variants.jsonl 2> $O/c18_variants.err
python - <<'PY'
import json
for l in open("gpurun_out/c18_variants.jsonl"):
    d = json.loads(l)
    print(d.get("spec"), d.get("error", "")[:200], d.get("parity_failures"), {k: (d[k]["ms_render"], d[k]["tail_ms"], d[k]["ms_render_share8"], d[k]["tail_ms_share8"]) for k in ("sponza", "dielectrics") if k in d})
PY
