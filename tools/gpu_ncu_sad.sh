#!/bin/bash
# ncu capture of the multi-reference SAD kernel (run after the same build passed its tests)
set -u
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
timeout 300 $CMD > gpurun_out/plain_sad.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:ime_sad_multi" -s 4 -c 1 \
    -f -o gpurun_out/prof_sad $CMD > gpurun_out/ncu_sad.log 2>&1
echo "ncu sad exit $?"
