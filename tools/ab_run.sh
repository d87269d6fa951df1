#!/bin/bash
# One single-GPU session: GPU test suite, then kernel A/B (tools/build_variants.sh) on the device-resident legs.
# usage (GPU box): bash tools/ab_run.sh "variant tags" "workloads"   -> gpurun_out/ab_<tag>_<workload>.json
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
tags=${1:-"default"}; wls=${2:-"p16 p8"}
if [ -z "$SKIP_TESTS" ]; then
  timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/ab_pytest.log 2>&1; echo "pytest rc $?"; tail -3 gpurun_out/ab_pytest.log
fi
for t in $tags; do
  lib=""; [ $t != default ] && lib=$PWD/speech_enhancement_b200/lib/variants/$t/libse_b200.so
  for w in $wls; do
    args="--profile-only $w"
    SE_B200_LIB=$lib timeout 300 python bench.py $args --steps 5 --warmup 3 > gpurun_out/ab_${t}_$w.json 2> gpurun_out/ab_${t}_$w.err
    python - "$t" "$w" <<'P'
import json, sys
t, w = sys.argv[1:3]
try:
    l = [x for x in open(f"gpurun_out/ab_{t}_{w}.json") if x.startswith("{")][-1]
    d = json.loads(l)
    r = d.get("roofline", {})
    print(t, w, "xRT %.4g" % d.get("xrt", 0), "ms/step %.3f" % d.get("ms_per_step", 0), "kernel_ms %.3f" % d.get("kernel_ms", 0))
except Exception as e:
    print(t, w, "FAILED", e)
P
  done
done
[ -n "$LATENCY" ] && python tests/tools/percall_latency.py > gpurun_out/ab_percall.md 2>&1 && cat gpurun_out/ab_percall.md
exit 0
