#!/bin/bash
# Round 2, thirteenth GPU call: texture wrap without the integer division — variants on four scenes, then the suite.
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
rm -f $O/full_size_reference.jsonl
export VARIANT_SHARE=8 VARIANT_STEPS=20 VARIANT_TIMEOUT=200
VARIANT_SCENES=sponza,dielectrics,sibenik,dynamic timeout 900 python tools/variant_bench.py "off,queue=off" "auto,queue=auto" "fast_auto,queue=auto,arith=fast" > $O/c13_variants.jsonl 2> $O/c13_variants.err
python - <<'PY'
import json
for l in open("gpurun_out/c13_variants.jsonl"):
    d = json.loads(l)
    print(d.get("spec"), d.get("error", "")[:200], d.get("parity_failures"), d.get("goldens_worst"), {k: (d[k]["ms_render"], d[k]["ms_fxaa"], d[k]["ms_render_share8"]) for k in ("sponza", "dielectrics", "sibenik", "dynamic") if k in d})
PY
timeout 900 python -m pytest tests -m gpu -x -q > $O/c13_gpu_tests.log 2>&1; echo "pytest rc $?" >> $O/c13_gpu_tests.log
tail -3 $O/c13_gpu_tests.log
