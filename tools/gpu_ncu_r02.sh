#!/bin/bash
# ncu evidence for round 2 (run under gpurun, one GPU): launch list + --set full captures of the dense SAD launch and of
# the per-PU decision kernel, on tools/replay_picture.py (see there why not on the encoder process)
set -u
cd "$(dirname "$0")/.."
OUT=gpurun_out; mkdir -p $OUT
CMD="python tools/replay_picture.py 8000"
$CMD > $OUT/r02_replay_plain.log 2>&1 || { echo "plain run failed"; tail -5 $OUT/r02_replay_plain.log; exit 1; }
cat $OUT/r02_replay_plain.log
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file $OUT/r02_launches.csv $CMD > $OUT/ncu_launches.log 2>&1
echo "launch list rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:ime_sad_multi_kernel -s 1 -c 1 -o $OUT/r02_prof_sad -f $CMD > $OUT/ncu_sad.log 2>&1
echo "sad capture rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:me_pu_kernel -s 1 -c 1 -o $OUT/r02_prof_me_pu -f $CMD > $OUT/ncu_me_pu.log 2>&1
echo "me_pu capture rc=$?"
ls -la $OUT/*.ncu-rep $OUT/r02_launches.csv
