#!/bin/bash
# End-to-end (host buffers) A/B of the SE_MEM_HOST pipeline options: se_ctx_set_pipeline_lead bits and chunk sizes.
cd "$(dirname "$0")/.."
for cfg in "--pipe-mode 0" "--pipe-mode 1" "--pipe-mode 2" "--pipe-mode 3" "--pipe-mode 3 --chunk-mb 256"; do
  python bench.py --no-cpu --no-extra --steps 5 --warmup 3 $cfg 2>/dev/null | python -c "
import sys,json
d=json.loads([l for l in sys.stdin if l.startswith('{')][-1])
print('cfg [$cfg] p16 e2e %.4g (%.1f ms)  p8 e2e %.4g (%.1f ms)  resynth e2e %.4g' % (d['e2e']['value'], d['e2e']['ms_per_step'], d['workloads']['p8']['e2e']['value'], d['workloads']['p8']['e2e']['ms_per_step'], d['workloads']['resynth']['e2e']['value']))"
done
