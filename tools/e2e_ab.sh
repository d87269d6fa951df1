#!/bin/bash
# End-to-end (host buffers) A/B of the SE_MEM_HOST pipeline: environment / bench options per run.
cd "$(dirname "$0")/.."
run() {  # label, env assignment, args
  env $2 python bench.py --no-cpu --no-extra --steps 5 --warmup 3 $3 2>/dev/null | python -c "
import sys,json
d=json.loads([l for l in sys.stdin if l.startswith('{')][-1])
print('[$1 $3] p16 e2e %.4g (%.1f ms)  p8 e2e %.4g (%.1f ms)  resynth e2e %.4g' % (d['e2e']['value'], d['e2e']['ms_per_step'], d['workloads']['p8']['e2e']['value'], d['workloads']['p8']['e2e']['ms_per_step'], d['workloads']['resynth']['e2e']['value']))"
}
run "default" A=1 ""
run "no lead chunk, no upload-ahead" A=1 "--pipe-mode 0"
run "256 MiB chunks" A=1 "--chunk-mb 256"
