#!/bin/bash
# Multi-GPU evidence in one `gpurun --gpus 8` call: host topology, cfg4 through the host work queue at 1/2/4/8 GPUs
# (one process), bench.py under torchrun at N = 4 and 8 (one process per GPU) with the H2D-only floor in the line.
# usage: tools/gpu_scale.sh <tag>
TAG=${1:-sc}; O=gpurun_out/$TAG; mkdir -p $O
nvidia-smi topo -m > $O/topo.txt 2>&1; (lscpu | grep -i -E "model name|socket|numa|^cpu\(s\)"; free -g | head -2) > $O/host.txt 2>&1
timeout 1200 python tools/cfg4_queue_bench.py --gpus 1,2,4,8 --pool 256 --batches-per-gpu 4 > $O/cfg4_queue.jsonl 2> $O/cfg4_queue.err
echo "cfg4 exit $?"; cut -c1-330 $O/cfg4_queue.jsonl
for N in 4 8; do
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29510 + N)) \
      bench.py --gpus $N --steps 10 --warmup 3 > $O/bench_n$N.json 2> $O/bench_n$N.err
  echo "bench N=$N exit $?"
  python -c "
import json; d=json.load(open('$O/bench_n$N.json'))
print('N', d['n_gpus'], 'value', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'e2e ms', round(d['e2e']['ms_per_step'],2), 'h2d floor ms', round(d['e2e']['h2d_floor_ms'],2), 'wav16', round(d['e2e_wav16']['value'],1))"
done
