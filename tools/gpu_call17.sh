#!/bin/bash
# Round 2, seventeenth GPU call: the FAST unit compiled with -use_fast_math — the suite (full-size FAST rows), variants, the FAST bench line.
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
rm -f $O/full_size_reference.jsonl
timeout 900 python -m pytest tests -m gpu -x -q > $O/c17_gpu_tests.log 2>&1; echo "pytest rc $?" >> $O/c17_gpu_tests.log
tail -3 $O/c17_gpu_tests.log
grep "arithmetic=fast" $O/full_size_reference.jsonl | cut -c1-300
export VARIANT_SHARE=8 VARIANT_STEPS=20 VARIANT_TIMEOUT=240
VARIANT_SCENES=sponza,dielectrics,sibenik,dynamic timeout 900 python tools/variant_bench.py "exact,queue=auto" "fast,queue=auto,arith=fast" > $O/c17_variants.jsonl 2> $O/c17_variants.err
python - <<'PY'
import json
for l in open("gpurun_out/c17_variants.jsonl"):
    d = json.loads(l)
    print(d.get("spec"), d.get("error", "")[:300], d.get("parity_failures"), d.get("goldens_worst"), {k: (d[k]["ms_render"], d[k]["ms_render_share8"]) for k in ("sponza", "dielectrics", "sibenik", "dynamic") if k in d})
PY
timeout 400 python bench.py --steps 20 --warmup 3 --arithmetic fast > $O/c17_bench_fast.json 2> $O/c17_bench_fast.err; echo "bench rc $?" >> $O/c17_bench_fast.err
tail -n 1 $O/c17_bench_fast.err; cut -c1-260 $O/c17_bench_fast.json
