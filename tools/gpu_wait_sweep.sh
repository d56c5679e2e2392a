#!/bin/bash
# how a CTU worker waits for its PU search x how many worker threads: encode fps of config 2 through kvazaar_cuda
# usage: tools/gpu_wait_sweep.sh FRAMES "owf list" "threads-multiplier list (x nproc / 4)" "wait modes"
set -u
cd "$(dirname "$0")/.."
OUT=gpurun_out; mkdir -p $OUT
N=${1:-33}
K=kvazaar_b200/host/_build/kvazaar_cuda
T=$(nproc)
C=/tmp/kvz_clips; mkdir -p $C
[ -f $C/c2w.yuv ] || python tools/make_clip.py $C/c2w.yuv 1920 1080 $N 2
echo "nproc $T"
for owf in ${2:-3}; do for q in ${3:-4 6 8}; do for wm in ${4:-sync spin yield}; do
  th=$((T * q / 4))
  r=$(env KVZ_CUDA_WAIT=$wm KVZ_CUDA_MODE=batch timeout 600 $K -i $C/c2w.yuv --input-res 1920x1080 -o $OUT/w.hevc \
    --preset medium --me full32 --gop 8 -r 4 --threads $th --owf $owf -n $N 2>&1 | grep -E " FPS" | tr -d '\n')
  echo "owf $owf threads $th wait $wm : $r  md5 $(md5sum < $OUT/w.hevc | cut -c1-8)"
done; done; done
rm -f $OUT/w.hevc
