#!/bin/bash
# One gpurun call: GPU parity tests, bench, microbench, ncu launch list, ncu --set full of the fbank kernel.
# usage: tools/gpu_check.sh <tag> [skip-ncu]
set -u
TAG=${1:-run}
O=gpurun_out/$TAG
mkdir -p $O
python -m pytest tests -m gpu -x -q > $O/gpu_tests.log 2>&1; echo "pytest exit $?" >> $O/gpu_tests.log
tail -3 $O/gpu_tests.log
python bench.py --steps 20 --warmup 3 > $O/bench_line.json 2> $O/bench.err; echo "bench exit $?"
cat $O/bench_line.json
python tools/microbench.py > $O/microbench.json 2> $O/microbench.err; echo "microbench exit $?"
cat $O/microbench.json
MODE=${2:-all}
if [ "$MODE" != "skip-ncu" ]; then
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $O/plain_bench.log 2>&1 &&
  ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file $O/bench_launches.csv \
      python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $O/ncu_bench.log 2>&1
  echo "ncu launches exit $?"
  python tools/microbench.py --only fbank_f64 --reps 1 > $O/plain_mb.log 2>&1 &&
  ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:FbankKernel.double \
      -s 3 -c 1 -o $O/fbank_f64 python tools/microbench.py --only fbank_f64 --reps 1 > $O/ncu_fbank.log 2>&1
  echo "ncu fbank exit $?"
  [ "$MODE" = "all" ] && python tools/microbench.py --only acoustic --reps 1 > $O/plain_mb2.log 2>&1 &&
  ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:AcousticTileKernel \
      -s 3 -c 1 -o $O/acoustic_tile python tools/microbench.py --only acoustic --reps 1 > $O/ncu_acoustic.log 2>&1
  echo "ncu acoustic exit $?"
fi
