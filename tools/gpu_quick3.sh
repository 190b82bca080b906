#!/bin/bash
# quick GPU iteration: all parity tests, microbench + phase trace of both tile kernels, short bench (e2e)
TAG=${1:-q}; O=gpurun_out/$TAG; mkdir -p $O
bash tools/gpu_quick2.sh $TAG
python bench.py --steps 10 --warmup 3 --no-cpu-baseline > $O/bench.json 2> $O/bench.err
python -c "
import json; d=json.load(open('$O/bench.json'))
print('value', round(d['value'],1), 'ms', round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value'],1), round(d['e2e']['ms_per_step'],2), 'wav16', round(d['e2e_wav16']['value'],1), round(d['e2e_wav16']['ms_per_step'],2))
print(d['config']['kernel_ms_per_step'])"
