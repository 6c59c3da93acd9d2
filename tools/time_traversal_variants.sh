#!/bin/bash
# Times the traversal variants prepared at the end of round 1 beside the shipped library,
# on the four scenes, and runs the parity tests that exercise traversal against each variant.
# Run from the repository root in the build container; results land in gpurun_out/.
set -e
make -C cpu-raytracer_b200 all variants
B=cpu-raytracer_b200/build
/usr/local/graft/bin/gpurun --timeout 420 -- "
VARIANT_SCENES=sponza,sibenik,dynamic,dielectrics timeout 200 python tools/variant_bench.py cpu-raytracer_b200/libb200rt.so $B/libb200rt_octant.so $B/libb200rt_prefetch.so $B/libb200rt_octant_prefetch.so > gpurun_out/variants_threaded.log 2>&1
for v in octant prefetch octant_prefetch; do
  echo \"== \$v\" >> gpurun_out/variants_threaded_tests.log
  B200RT_LIB=\$PWD/$B/libb200rt_\$v.so timeout 90 python -m pytest tests/test_config_switches.py tests/test_device_edges.py tests/test_device_parity.py -m gpu -q 2>&1 | tail -3 >> gpurun_out/variants_threaded_tests.log
done
cat gpurun_out/variants_threaded.log gpurun_out/variants_threaded_tests.log"
