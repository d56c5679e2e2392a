#!/bin/bash
# config 2 (1080p, medium, full32, gop 8, 4 refs) through kvazaar_cuda in batch mode: timing + per-call statistics
# usage: tools/gpu_e2e_c2.sh FRAMES [extra env assignments...]
set -u
cd "$(dirname "$0")/.."
OUT=gpurun_out; mkdir -p $OUT
N=${1:-9}; shift || true
K=${KBIN:-kvazaar_b200/host/_build/kvazaar_cuda}
T=$(nproc)
C=/tmp/kvz_clips; mkdir -p $C
[ -f $C/c2.yuv ] || python tools/make_clip.py $C/c2.yuv 1920 1080 17 2
env "$@" KVZ_CUDA_STATS=${STATS:-1} KVZ_CUDA_MODE=batch timeout 900 $K -i $C/c2.yuv --input-res 1920x1080 -o $OUT/c2.hevc \
  --preset medium --me full32 --gop 8 -r 4 --threads ${THREADS:-$T} --owf ${OWF:-0} -n $N 2>&1 | grep -E "FPS|kvz_|wall"
md5sum $OUT/c2.hevc; rm -f $OUT/c2.hevc
