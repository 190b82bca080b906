#!/bin/bash
# quick GPU iteration: parity tests, microbench of one kernel family, its phase trace
# usage: tools/gpu_quick.sh <tag> <fbank_f64|acoustic|...> [pytest -k expr]
TAG=${1:-q}; WHAT=${2:-acoustic}; K=${3:-}
O=gpurun_out/$TAG; mkdir -p $O
if [ -n "$K" ]; then python -m pytest tests -m gpu -x -q -k "$K" > $O/tests.log 2>&1; else python -m pytest tests -m gpu -x -q > $O/tests.log 2>&1; fi
tail -3 $O/tests.log
python tools/microbench.py --only $WHAT > $O/mb.json 2> $O/mb.err; cat $O/mb.json
AFE_PHASE_TRACE=1 python tools/microbench.py --only $WHAT --reps 1 > /dev/null 2> $O/trace.err; grep "afe trace" $O/trace.err | tail -9
