#!/bin/bash
# Round 2, twenty-seventh GPU call: which child the exporting thread keeps (reflection / refraction); then suite + stress with the shipped setting.
cd "$(dirname "$0")/.."
O=gpurun_out
export VARIANT_SHARE=8 VARIANT_STEPS=15 VARIANT_TIMEOUT=240
VARIANT_SCENES=sponza,dielectrics,dynamic timeout 1200 python tools/variant_bench.py "off,queue=off" "all_keep_reflection,queue=all,keep=0" "all_keep_refraction,queue=all,keep=1" "glass_keep_reflection,queue=glass,keep=0" "glass_keep_refraction,queue=glass,keep=1" > $O/c27_variants.jsonl 2> $O/c27_variants.err
python - <<'PY'
import json
for l in open("gpurun_out/c27_variants.jsonl"):
    d = json.loads(l)
    print(d.get("spec"), d.get("error", "")[:200], d.get("parity_failures"), {k: (d[k]["ms_render"], d[k]["tail_ms"], d[k]["ms_render_share8"], d[k]["tail_ms_share8"]) for k in ("sponza", "dielectrics", "dynamic") if k in d})
PY
rm -f $O/full_size_reference.jsonl
timeout 900 python -m pytest tests -m gpu -x -q > $O/c27_gpu_tests.log 2>&1; echo "pytest rc $?" >> $O/c27_gpu_tests.log
tail -3 $O/c27_gpu_tests.log
STRESS_TIMEOUT=200 timeout 1300 python tools/queue_stress.py 1000 > $O/c27_queue_stress.jsonl 2> $O/c27_queue_stress.err; cat $O/c27_queue_stress.jsonl
B200RT_QUEUE_KEEP_REFRACTION=1 STRESS_TIMEOUT=200 timeout 1300 python tools/queue_stress.py 500 > $O/c27_queue_stress_keep_refraction.jsonl 2>> $O/c27_queue_stress.err; cat $O/c27_queue_stress_keep_refraction.jsonl
