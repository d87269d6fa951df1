#!/bin/bash
# End-to-end (host buffers) A/B of the SE_MEM_HOST pipeline: library variant (SE_B200_LIB) x chunk size.
cd "$(dirname "$0")/.."
run() {  # label, lib, args
  SE_B200_LIB=$2 python bench.py --no-cpu --no-extra --steps 5 --warmup 3 $3 2>/dev/null | python -c "
import sys,json
d=json.loads([l for l in sys.stdin if l.startswith('{')][-1])
print('[$1 $3] p16 e2e %.4g (%.1f ms)  p8 e2e %.4g (%.1f ms)  resynth e2e %.4g' % (d['e2e']['value'], d['e2e']['ms_per_step'], d['workloads']['p8']['e2e']['value'], d['workloads']['p8']['e2e']['ms_per_step'], d['workloads']['resynth']['e2e']['value']))"
}
L6=$PWD/speech_enhancement_b200/lib/variants/l6/libse_b200.so
run "3 lanes" "" ""
run "6 lanes" $L6 ""
run "6 lanes" $L6 "--chunk-mb 256"
run "6 lanes" $L6 "--chunk-mb 128 --chunk-utts 512"
