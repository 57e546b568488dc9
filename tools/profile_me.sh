#!/bin/sh
# ncu --set full capture of the motion-search kernels of one bench.py ME step, exported here
# (the .ncu-rep is too big to travel): raw metrics + per-SASS-line counters with CUDA source lines.
set -x
ncu --set full --clock-control none --import-source on -k regex:pel_kernel --launch-skip 6 -c 2 \
  -o gpurun_out/me_full python bench.py --sizes 0 --steps 2 --warmup 3 --no-cpu --no-extras --e2e-steps 0 \
  > gpurun_out/ncu_me.log 2>&1
ncu -i gpurun_out/me_full.ncu-rep --page raw --csv > gpurun_out/me_raw.csv
ncu -i gpurun_out/me_full.ncu-rep --page source --csv --print-source cuda,sass > gpurun_out/me_src_cuda.csv 2> gpurun_out/me_src_cuda.err
ncu -i gpurun_out/me_full.ncu-rep --page source --csv --print-source sass > gpurun_out/me_src.csv
rm -f gpurun_out/me_full.ncu-rep
ls -la gpurun_out | tail -8
