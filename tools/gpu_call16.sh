#!/bin/bash
# Round 2, sixteenth GPU call: the FAST unit with approximate division / square root, and with all of -use_fast_math.
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
export VARIANT_SHARE=0 VARIANT_STEPS=20 VARIANT_TIMEOUT=240
B=cpu-raytracer_b200/build
VARIANT_SCENES=sponza,dielectrics,sibenik,dynamic timeout 900 python tools/variant_bench.py "exact,queue=auto" "fast,queue=auto,arith=fast" "fast_fastdiv,lib=$B/libb200rt_fastdiv.so,queue=auto,arith=fast" "fast_fastmath,lib=$B/libb200rt_fastmath.so,queue=auto,arith=fast" > $O/c16_variants.jsonl 2> $O/c16_variants.err
python - <<'PY'
import json
for l in open("gpurun_out/c16_variants.jsonl"):
    d = json.loads(l)
    print(d.get("spec"), d.get("error", "")[:300], d.get("parity_failures"), d.get("goldens_worst"), {k: d[k]["ms_render"] for k in ("sponza", "dielectrics", "sibenik", "dynamic") if k in d})
PY
