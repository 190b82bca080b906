#!/bin/bash
# One quick GPU iteration (a single gpurun call): parity tests, microbench + sub-phase trace of one kernel family.
# usage: tools/gpu_iter.sh <tag> <fbank_f64|acoustic|all> [pytest -k expr]      (every step runs under its own timeout)
TAG=${1:-q}; WHAT=${2:-all}; K=${3:-}
O=gpurun_out/$TAG; mkdir -p $O
if [ -n "$K" ]; then timeout 900 python -m pytest tests -m gpu -x -q -k "$K" > $O/tests.log 2>&1; else timeout 900 python -m pytest tests -m gpu -x -q > $O/tests.log 2>&1; fi
echo "pytest exit $?"; tail -3 $O/tests.log
FAMS="$WHAT"; [ "$WHAT" = "all" ] && FAMS="fbank_f64 acoustic"
for W in $FAMS; do
  timeout 300 python tools/microbench.py --only $W > $O/mb_$W.json 2> $O/mb_$W.err; echo "microbench $W exit $?"
  grep -A5 '"kernels"' $O/mb_$W.json | tr -d '\n '; echo
  AFE_PHASE_TRACE=1 timeout 300 python tools/microbench.py --only $W --reps 1 > /dev/null 2> $O/trace_$W.err
  grep "afe trace" $O/trace_$W.err | tail -9 | cut -c1-200
done
