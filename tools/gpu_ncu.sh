#!/bin/bash
# ncu --set full capture of one kernel of tools/microbench.py (a single gpurun call).
# usage: tools/gpu_ncu.sh <tag> <microbench family: fbank_f64|acoustic> <kernel regex> [name]
TAG=${1:-n}; FAM=${2:-fbank_f64}; RX=${3:-fbank_pipe_kernel}; NAME=${4:-$FAM}
O=gpurun_out/$TAG; mkdir -p $O
timeout 300 python tools/microbench.py --only $FAM --reps 1 > $O/plain_mb_$NAME.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:$RX \
    -s 3 -c 1 -o $O/$NAME python tools/microbench.py --only $FAM --reps 1 > $O/ncu_$NAME.log 2>&1
echo "ncu $NAME exit $?"; tail -3 $O/ncu_$NAME.log
