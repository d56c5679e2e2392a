#!/bin/bash
# tools/gpu_check.sh -- what one gpurun call does: GPU parity tests, a short bench, the
# per-launch ncu list and one full capture of the dominant kernel.  Usage (from the repo root):
#   gpurun --timeout 1500 -- 'bash tools/gpu_check.sh [notests] [noncu]'
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,power.limit --format=csv > gpurun_out/gpu.txt 2>&1
nproc > gpurun_out/nproc.txt
if [[ " $* " != *" notests "* ]]; then
  timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1
  echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
  tail -5 gpurun_out/pytest_gpu.log
fi
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err
echo "bench exit $?"; cat gpurun_out/bench.json; tail -3 gpurun_out/bench.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
echo "bench ref exit $?"; cat gpurun_out/bench_ref.json
if [[ " $* " != *" noncu "* ]]; then
  CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
  timeout 300 $CMD > gpurun_out/plain.log 2>&1 &&
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
      --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
  echo "ncu list exit $?"
  timeout 300 $CMD > gpurun_out/plain2.log 2>&1 &&
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:ime_sad_multi -s 4 -c 1 \
      -f -o gpurun_out/prof_sad $CMD > gpurun_out/ncu_full.log 2>&1
  echo "ncu full exit $?"
fi
