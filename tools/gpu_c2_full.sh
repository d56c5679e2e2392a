#!/bin/bash
# BASELINE config 2 in full: 1920x1080 8-bit, 64 pictures, --preset medium --me full32 --gop 8 -r 4 (owf 3 = what --owf auto
# gives on 16 threads), the unmodified reference CLI and kvazaar_cuda on the same box: wall time of each, md5 of both bitstreams
set -u
cd "$(dirname "$0")/.."
OUT=gpurun_out; mkdir -p $OUT
T=$(nproc)
C=/tmp/kvz_clips; mkdir -p $C
[ -f $C/c2_64.yuv ] || python tools/make_clip.py $C/c2_64.yuv 1920 1080 64 2
OPTS="--input-res 1920x1080 --preset medium --me full32 --gop 8 -r 4 --owf 3 --threads $T -n 64"
echo "host threads $T; options: $OPTS"
echo "== kvazaar_cuda (KVZ_CUDA_MODE=batch)"
KVZ_CUDA_MODE=batch timeout -s KILL 300 kvazaar_b200/host/_build/kvazaar_cuda -i $C/c2_64.yuv -o $OUT/c2_64_cuda.hevc $OPTS 2>&1 | grep -E "FPS|wall|Bits|PSNR" | tail -4
md5sum $OUT/c2_64_cuda.hevc
echo "== reference (oracle/_ref/kvazaar, unmodified)"
timeout -s KILL 600 oracle/_ref/kvazaar -i $C/c2_64.yuv -o $OUT/c2_64_ref.hevc $OPTS 2>&1 | grep -E "FPS|wall|Bits|PSNR" | tail -4
md5sum $OUT/c2_64_ref.hevc
cmp $OUT/c2_64_cuda.hevc $OUT/c2_64_ref.hevc && echo "bitstreams identical ($(stat -c %s $OUT/c2_64_ref.hevc) bytes)"
rm -f $OUT/c2_64_cuda.hevc $OUT/c2_64_ref.hevc
