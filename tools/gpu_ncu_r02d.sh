#!/bin/bash
# ncu evidence for the TMA staging of the dense SAD launch (run under gpurun, one GPU), on tools/replay_picture.py:
# plain runs with TMA and with cp.async staging, then one --set full capture of ime_sad_multi_tma_kernel
set -u
cd "$(dirname "$0")/.."
OUT=gpurun_out; mkdir -p $OUT
CMD="python tools/replay_picture.py 8000"
for t in 1 0 1 0; do echo "== KVZ_CUDA_SAD_TMA=$t"; KVZ_CUDA_SAD_TMA=$t timeout -s KILL 120 $CMD 2>&1 | tail -1; done | tee $OUT/r02d_replay_tma_vs_cpasync.txt
timeout -s KILL 500 ncu --set full --clock-control none --import-source on -k regex:ime_sad_multi_tma_kernel -s 1 -c 1 -o $OUT/r02d_prof_sad_tma -f $CMD > $OUT/ncu_sad_tma.log 2>&1
echo "sad tma capture rc=$?"
ls -la $OUT/r02d_prof_sad_tma.ncu-rep
