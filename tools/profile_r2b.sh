#!/bin/sh
# ncu --set full captures of the pixel-input transform kernels (config[3]) and the inverse-transform
# kernels; reports exported to CSV here (too big to travel).
set -x
B="python bench.py --sizes 0 --steps 2 --warmup 1 --no-cpu --e2e-steps 0 --config5-frames 4"
ncu --set full --clock-control none --import-source on -k 'regex:txq.*, 2>' --launch-skip 19 -c 19 \
  -o gpurun_out/hbd_full $B > gpurun_out/ncu_hbd.log 2>&1
ncu -i gpurun_out/hbd_full.ncu-rep --page raw --csv > gpurun_out/hbd_raw.csv
rm -f gpurun_out/hbd_full.ncu-rep
ncu --set full --clock-control none --import-source on -k regex:itx_kernel --launch-skip 3 -c 73 \
  -o gpurun_out/itx_full $B > gpurun_out/ncu_itx.log 2>&1
ncu -i gpurun_out/itx_full.ncu-rep --page raw --csv > gpurun_out/itx_raw.csv
rm -f gpurun_out/itx_full.ncu-rep
ls -la gpurun_out
