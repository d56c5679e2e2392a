#!/bin/bash
# development aid (GPU box): short bench, prints step time, SAD time and roofline fraction
python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_q.json 2> gpurun_out/bench_q.err
python -c "
import json; d=json.load(open('gpurun_out/bench_q.json')); print('value', round(d['value'],1), 'ms', round(d['ms_per_step'],3), 'frac', round(d['roofline']['frac'],4), d['config']['kernel_ms'])"
tail -3 gpurun_out/bench_q.err
