#!/bin/bash
# 8-GPU box: end-to-end A/B of the upload-ahead pipeline (se_ctx_set_pipeline_lead bit 2) at N = 8 and N = 4.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
port=29610
for cfg in "8 1" "8 3" "4 1" "4 3" "2 3"; do
  set -- $cfg
  port=$((port+1))
  timeout 400 $TR --nproc-per-node $1 --master-port $port bench.py --gpus $1 --pipe-mode $2 --no-cpu --no-extra --steps 4 --warmup 3 \
      2> gpurun_out/r02_scale_ab_n$1_m$2.err | grep '^{' > gpurun_out/r02_scale_ab_n$1_m$2.json
  python - "$1" "$2" <<'P'
import json, sys
n, m = sys.argv[1:3]
try:
    d = json.loads(open(f"gpurun_out/r02_scale_ab_n{n}_m{m}.json").read().strip().splitlines()[-1])
    e = d["e2e"]
    print("N=%s mode %s: p16 e2e %.4g (%.1f ms, eff %s)  p8 e2e %.4g  resynth e2e %.4g" % (n, m, e["value"], e["ms_per_step"], e.get("e2e_efficiency"),
          d["workloads"]["p8"]["e2e"]["value"], d["workloads"]["resynth"]["e2e"]["value"]))
except Exception as ex:
    print("N=%s mode %s FAILED %r" % (n, m, ex))
P
done
