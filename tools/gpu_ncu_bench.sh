#!/bin/bash
# ncu evidence of the BENCH step itself (cfg2, one gpurun call): launch list of the whole run and one `--set full`
# capture of the fbank launch of the first TIMED device-resident step (one launch per step covers the three streams; launches 0-2 are the warm-up steps).
# usage: tools/gpu_ncu_bench.sh <tag>
TAG=${1:-nb}; O=gpurun_out/$TAG; mkdir -p $O
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
timeout 600 $CMD > $O/plain_bench.json 2> $O/plain_bench.err &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file $O/bench_launches.csv \
    $CMD > $O/ncu_launches.log 2>&1
echo "ncu launch list exit $?"
timeout 900 ncu --set full --clock-control none --import-source on --kernel-name-base demangled \
    -k regex:FbankKernel.double -s 3 -c 1 -o $O/fbank_cfg2 $CMD > $O/ncu_fbank_cfg2.log 2>&1
echo "ncu fbank cfg2 exit $?"; tail -2 $O/ncu_fbank_cfg2.log
