#!/bin/bash
# Round 2, tenth GPU call: the L2 warm-up prologue of the render kernels, on and off, flushed and warm L2, whole frame and 8-GPU share.
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
export VARIANT_SHARE=8 VARIANT_STEPS=20 VARIANT_TIMEOUT=200
VARIANT_SCENES=sponza,dielectrics,sibenik,dynamic timeout 900 python tools/variant_bench.py "warm_off,queue=off,warm=0" "warm_on,queue=off,warm=1" "warm_off_l2warm,queue=off,warm=0,flush=0" "warm_on_l2warm,queue=off,warm=1,flush=0" \
   "auto_warm_off,queue=auto,warm=0" "auto_warm_on,queue=auto,warm=1" "fast_warm_on,queue=auto,arith=fast,warm=1" > $O/c10_variants.jsonl 2> $O/c10_variants.err
python - <<'PY'
import json
for l in open("gpurun_out/c10_variants.jsonl"):
    d = json.loads(l)
    print(d.get("spec"), d.get("error", "")[:200], {k: (d[k]["ms_render"], d[k]["tail_ms"], d[k]["ms_render_share8"], d[k]["tail_ms_share8"]) for k in ("sponza", "dielectrics", "sibenik", "dynamic") if k in d})
PY
timeout 900 python -m pytest tests -m gpu -x -q > $O/c10_gpu_tests.log 2>&1; echo "pytest rc $?" >> $O/c10_gpu_tests.log
tail -3 $O/c10_gpu_tests.log
