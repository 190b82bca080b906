#!/bin/bash
# quick GPU iteration over BOTH tile kernels: all parity tests, microbench + phase trace of fbank_f64 and acoustic
TAG=${1:-q}; O=gpurun_out/$TAG; mkdir -p $O
python -m pytest tests -m gpu -x -q > $O/tests.log 2>&1; tail -2 $O/tests.log
for W in fbank_f64 acoustic; do
  python tools/microbench.py --only $W > $O/mb_$W.json 2> $O/mb_$W.err; grep -A4 '"kernels"' $O/mb_$W.json | tr -d '\n'; echo
  AFE_PHASE_TRACE=1 python tools/microbench.py --only $W --reps 1 > /dev/null 2> $O/trace_$W.err; grep "afe trace" $O/trace_$W.err | tail -9 | cut -c1-90
done
