#!/bin/bash
# Round 2, twenty-second GPU call: how full are the warps that walk queue items (trace variant).
cd "$(dirname "$0")/.."
O=gpurun_out
export B200RT_LIB=$PWD/cpu-raytracer_b200/build/libb200rt_trace.so
for spec in "sponza 1 all" "sponza 8 all" "dielectrics 1 all" "dielectrics 8 all"; do timeout 200 python tools/queue_trace.py $spec 2>&1 | grep -v "^$"; done > $O/c22_queue_trace.txt 2>&1
cat $O/c22_queue_trace.txt
