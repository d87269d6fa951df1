#!/bin/bash
# One 8-GPU session: the N-rank host<->device copy probe (both GPU maps) and bench.py at N = 2, 4, 8.
# Usage (on the GPU box): bash tools/scale_run.sh   -> gpurun_out/r02_pcie_n*.json, r02_bench_n*.json
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
nvidia-smi topo -m > gpurun_out/r02_topo.txt 2>&1
nproc > gpurun_out/r02_nproc.txt
port=29510
for n in 1 2 4 8; do
  for map in stride dense; do
    if [ $n = 1 ] && [ $map = dense ]; then continue; fi
    if [ $n = 8 ] && [ $map = dense ]; then continue; fi
    port=$((port+1))
    timeout 300 $TR --nproc-per-node $n --master-port $port tools/pcie_probe.py --gpu-map $map 2>/dev/null | grep '^{' > gpurun_out/r02_pcie_n${n}_${map}.json
  done
done
for cfg in "2 stride" "2 dense" "4 stride" "8 stride"; do
  set -- $cfg
  port=$((port+1))
  timeout 900 $TR --nproc-per-node $1 --master-port $port bench.py --gpus $1 --gpu-map $2 --steps 5 --warmup 3 \
      > gpurun_out/r02_bench_n$1_$2.json 2> gpurun_out/r02_bench_n$1_$2.err
  echo "bench N=$1 $2 rc $?"
done
