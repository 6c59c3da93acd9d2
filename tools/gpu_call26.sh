#!/bin/bash
# Round 2, twenty-sixth GPU call: the exporting thread goes on with one child itself; only a fan-out queues an item.
cd "$(dirname "$0")/.."
O=gpurun_out
export VARIANT_SHARE=8 VARIANT_STEPS=15 VARIANT_TIMEOUT=240
VARIANT_SCENES=sponza,dielectrics,dynamic timeout 1200 python tools/variant_bench.py "off,queue=off" "auto,queue=auto" "all,queue=all" "glass,queue=glass" > $O/c26_variants.jsonl 2> $O/c26_variants.err
python - <<'PY'
import json
for l in open("gpurun_out/c26_variants.jsonl"):
    d = json.loads(l)
    print(d.get("spec"), d.get("error", "")[:200], d.get("parity_failures"), {k: (d[k]["ms_render"], d[k]["tail_ms"], d[k]["ms_render_share8"], d[k]["tail_ms_share8"]) for k in ("sponza", "dielectrics", "dynamic") if k in d})
PY
export B200RT_LIB=$PWD/cpu-raytracer_b200/build/libb200rt_trace.so
for spec in "sponza 8 all" "dielectrics 8 all"; do timeout 200 python tools/queue_trace.py $spec 2>&1 | grep -v "^$"; done > $O/c26_queue_trace.txt 2>&1
cat $O/c26_queue_trace.txt
