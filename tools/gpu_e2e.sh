#!/bin/bash
# End-to-end runs of the reference encoder with the CUDA path bound in (kvazaar_b200/host/_build/kvazaar_cuda)
# next to the unmodified reference (oracle/_ref/kvazaar) on the same clips and options; prints md5s and timings.
# usage: tools/gpu_e2e.sh [quick]
set -u
cd "$(dirname "$0")/.."
OUT=gpurun_out; mkdir -p $OUT
K=kvazaar_b200/host/_build/kvazaar_cuda
R=oracle/_ref/kvazaar
T=$(nproc)
C=/tmp/kvz_clips; mkdir -p $C
python tools/make_clip.py $C/c1.yuv 416 240 16 1
python tools/make_clip.py $C/c2s.yuv 416 240 17 2
python tools/make_clip.py $C/c2.yuv 1920 1080 17 2
run() {  # name, binary, mode, clip, res, options...
  local name=$1 bin=$2 mode=$3 clip=$4 res=$5; shift 5
  local t0=$(date +%s.%N)
  KVZ_CUDA_STATS=1 KVZ_CUDA_MODE=$mode timeout 900 $bin -i $clip --input-res $res -o $OUT/$name.hevc "$@" > $OUT/$name.log 2>&1
  local rc=$?
  local t1=$(date +%s.%N)
  echo "$name rc=$rc wall=$(echo "$t1 - $t0" | bc) md5=$(md5sum < $OUT/$name.hevc | cut -c1-32) $(grep -E 'FPS' $OUT/$name.log | tr -s ' ')"
  grep -E "kvz_glue|kvz_cuda_enc" $OUT/$name.log
  rm -f $OUT/$name.hevc
}
echo "host threads: $T"
C1="--preset ultrafast -q 32 --gop lp-g4d3t1 --threads $T --owf 0"
run c1_ref   $R off     $C/c1.yuv 416x240 $C1
run c1_batch $K batch   $C/c1.yuv 416x240 $C1
run c1_percall4 $K percall $C/c1.yuv 416x240 $C1 -n 4
run c1_ref4  $R off     $C/c1.yuv 416x240 $C1 -n 4
C2S="--preset medium --me full32 --gop 8 -r 4 --threads $T --owf 0"
run c2s_ref   $R off   $C/c2s.yuv 416x240 $C2S
run c2s_batch $K batch $C/c2s.yuv 416x240 $C2S
if [ "${1:-}" != quick ]; then
  run c2_batch9 $K batch $C/c2.yuv 1920x1080 $C2S -n 9
  run c2_ref_hexbs9 $R off $C/c2.yuv 1920x1080 --preset medium --gop 8 -r 4 --threads $T --owf 0 -n 9
  run c2_ref3 $R off $C/c2.yuv 1920x1080 $C2S -n 3
  run c2_batch3 $K batch $C/c2.yuv 1920x1080 $C2S -n 3
fi
