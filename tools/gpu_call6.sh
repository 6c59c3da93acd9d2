#!/bin/bash
# Round 2, sixth GPU call: the entry point that allocates nothing — its test alone in a fresh process, then the whole gpu suite with it last.
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > $O/c6_no_alloc_alone.log 2>&1; echo "pytest rc $?" >> $O/c6_no_alloc_alone.log
tail -5 $O/c6_no_alloc_alone.log
timeout 1500 python -m pytest tests -m gpu -x -q > $O/c6_gpu_tests.log 2>&1; echo "pytest rc $?" >> $O/c6_gpu_tests.log
tail -5 $O/c6_gpu_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/c6_smoke.log 2>&1; echo "smoke rc $?" >> $O/c6_smoke.log; tail -3 $O/c6_smoke.log
