#!/bin/bash
# Round 2, twelfth GPU call: one 64-bit reservation per export group — suite, stress, variants.
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
rm -f $O/full_size_reference.jsonl
export VARIANT_SHARE=8 VARIANT_STEPS=15 VARIANT_TIMEOUT=200
timeout 900 python tools/variant_bench.py "off,queue=off" "auto,queue=auto" "queue_all,queue=all" "queue_glass,queue=glass" "fast_auto,queue=auto,arith=fast" > $O/c12_variants.jsonl 2> $O/c12_variants.err
VARIANT_SCENES=sibenik,dynamic timeout 400 python tools/variant_bench.py "off,queue=off" "queue_all,queue=all" >> $O/c12_variants.jsonl 2>> $O/c12_variants.err
python - <<'PY'
import json
for l in open("gpurun_out/c12_variants.jsonl"):
    d = json.loads(l)
    print(d.get("spec"), d.get("error", "")[:200], d.get("parity_failures"), {k: (d[k]["ms_render"], d[k]["tail_ms"], d[k]["ms_render_share8"], d[k]["tail_ms_share8"]) for k in ("sponza", "dielectrics", "sibenik", "dynamic") if k in d})
PY
STRESS_TIMEOUT=200 timeout 1300 python tools/queue_stress.py 1000 > $O/c12_queue_stress.jsonl 2> $O/c12_queue_stress.err
cat $O/c12_queue_stress.jsonl
timeout 900 python -m pytest tests -m gpu -x -q > $O/c12_gpu_tests.log 2>&1; echo "pytest rc $?" >> $O/c12_gpu_tests.log
tail -3 $O/c12_gpu_tests.log
