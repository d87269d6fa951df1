#!/bin/bash
# ncu --set full capture of one bench leg's kernel (after the same command ran clean without ncu), exported to CSV.
# usage (GPU box): bash tools/ncu_run.sh <profile-only workload> <kernel regex> <out tag> [skip launches]
cd "$(dirname "$0")/.."
w=$1; k=$2; tag=$3; skip=${4:-3}
mkdir -p gpurun_out
python bench.py --profile-only $w --steps 1 --warmup 3 > gpurun_out/${tag}_plain.json 2> gpurun_out/${tag}_plain.err || { echo "plain run failed"; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:$k -s $skip -c 1 -f -o gpurun_out/$tag \
    python bench.py --profile-only $w --steps 1 --warmup 3 > gpurun_out/${tag}_ncu.log 2>&1 || { tail -5 gpurun_out/${tag}_ncu.log; exit 1; }
ncu -i gpurun_out/$tag.ncu-rep --page raw --csv > gpurun_out/${tag}_raw.csv
ncu -i gpurun_out/$tag.ncu-rep --page source --csv --print-source sass,cuda > gpurun_out/${tag}_src.csv
python tools/ncu_hot_lines.py gpurun_out/${tag}_src.csv 30 > gpurun_out/${tag}_hot.txt 2>&1
ls -la gpurun_out/$tag.ncu-rep
