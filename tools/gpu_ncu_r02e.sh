#!/bin/bash
# ncu --set full of the block kernels (rows A13-A17: transforms, quantiser, fused residual, RDOQ, intra costs, CABAC bit
# cost), one launch per stage of tools/time_stages.py at 1080p frame scale (run under gpurun, one GPU)
set -u
cd "$(dirname "$0")/.."
OUT=gpurun_out; mkdir -p $OUT
timeout -s KILL 300 python tools/time_stages.py > $OUT/r02e_stage_times.json 2> $OUT/stages.err || { echo "plain run failed"; tail -5 $OUT/stages.err; exit 1; }
echo "plain run ok"
KVZ_STAGES_ONCE=1 timeout -s KILL 900 ncu --set full --clock-control none \
  -k regex:'transform_small|transform_rows|quant_groups|quant_batch|quantize_residual|rdoq_kernel|intra_costs|coeff_bits' -c 40 \
  -o $OUT/r02e_prof_blocks -f python tools/time_stages.py > $OUT/ncu_blocks.log 2>&1
echo "capture rc=$?"
# the report of 30-40 full captures exceeds what gpurun copies back: keep the raw page as CSV, drop the report when large
timeout 300 ncu -i $OUT/r02e_prof_blocks.ncu-rep --page raw --csv > $OUT/r02e_prof_blocks_raw.csv 2> /dev/null
ls -la $OUT/r02e_prof_blocks.ncu-rep $OUT/r02e_prof_blocks_raw.csv
[ $(stat -c %s $OUT/r02e_prof_blocks.ncu-rep) -gt 30000000 ] && rm -f $OUT/r02e_prof_blocks.ncu-rep
