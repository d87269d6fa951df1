#!/bin/bash
# One 8-GPU session: the N-rank host<->device copy probe (both GPU maps) and bench.py at N = 2, 4, 8.
# Usage (on the GPU box): bash tools/scale_run.sh   -> gpurun_out/r02_pcie_n*.json, r02_bench_n*.json
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
nvidia-smi topo -m > gpurun_out/r02_topo.txt 2>&1
nproc > gpurun_out/r02_nproc.txt
(numactl -H || lscpu | grep -i numa) > gpurun_out/r02_numa.txt 2>&1
port=29510
for cfg in "2 stride" "2 dense" "4 stride" "4 dense" "8 stride"; do
  set -- $cfg
  port=$((port+1))
  timeout 200 $TR --nproc-per-node $1 --master-port $port tools/pcie_probe.py --gpu-map $2 --reps 3 2>/dev/null | grep '^{' > gpurun_out/r02_pcie_n$1_$2.json
done
for cfg in "2 stride --no-extra" "2 dense --no-extra" "4 stride --no-extra" "8 stride --no-cpu"; do
  set -- $cfg
  port=$((port+1))
  timeout 600 $TR --nproc-per-node $1 --master-port $port bench.py --gpus $1 --gpu-map $2 $3 --steps 5 --warmup 3 \
      > gpurun_out/r02_bench_n$1_$2.json 2> gpurun_out/r02_bench_n$1_$2.err
  echo "bench N=$1 $2 rc $?"
done
