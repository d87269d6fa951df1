cd /root/repo
python tests/tools/pipe_profile.py 3 || exit 1
ncu --set full --clock-control none --import-source on -k regex:ns_pipe_kernel -s 1 -c 1 -f -o gpurun_out/r02_pipe8 python tests/tools/pipe_profile.py 1 > gpurun_out/r02_pipe8_ncu.log 2>&1
ncu -i gpurun_out/r02_pipe8.ncu-rep --page source --csv --print-source sass,cuda > gpurun_out/r02_pipe8_src.csv
ncu -i gpurun_out/r02_pipe8.ncu-rep --page raw --csv > gpurun_out/r02_pipe8_raw.csv
python tools/ncu_hot_lines.py gpurun_out/r02_pipe8_src.csv 40
