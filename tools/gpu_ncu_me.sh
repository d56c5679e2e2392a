#!/bin/bash
# ncu capture of the decision kernels (run after tools/gpu_check.sh passed with the same build)
set -u
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
timeout 300 $CMD > gpurun_out/plain_me.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:me_scan|me_refine" -s 8 -c 4 \
    -f -o gpurun_out/prof_me $CMD > gpurun_out/ncu_me.log 2>&1
echo "ncu me exit $?"
