#!/bin/bash
# Round 2, fifteenth GPU call: EWA weight table in global memory instead of the constant bank.
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
export VARIANT_SHARE=8 VARIANT_STEPS=20 VARIANT_TIMEOUT=200
VARIANT_SCENES=sponza,dielectrics timeout 600 python tools/variant_bench.py "off,queue=off" "auto,queue=auto" "fast_auto,queue=auto,arith=fast" > $O/c15_variants.jsonl 2> $O/c15_variants.err
python - <<'PY'
import json
for l in open("gpurun_out/c15_variants.jsonl"):
    d = json.loads(l)
    print(d.get("spec"), d.get("error", "")[:200], d.get("parity_failures"), d.get("goldens_worst"), {k: (d[k]["ms_render"], d[k]["ms_fxaa"], d[k]["ms_render_share8"]) for k in ("sponza", "dielectrics") if k in d})
PY
timeout 600 python -m pytest tests/test_device_parity.py tests/test_config_switches.py tests/test_full_size_reference.py -m gpu -x -q -k "not stock" 2>&1 | tail -2
