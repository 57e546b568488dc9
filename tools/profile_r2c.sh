#!/bin/sh
# ncu --set full capture of the pixel-input transform kernels (config[3]: 4K 10-bit, uint16 planes).
set -x
B="python bench.py --only hbd4k --steps 2 --warmup 1"
$B > gpurun_out/hbd4k_plain.json 2> gpurun_out/hbd4k_plain.err || exit 1
ncu --set full --clock-control none --import-source on -k regex:txq --launch-skip 19 -c 19 \
  -o gpurun_out/hbd_full $B > gpurun_out/ncu_hbd.log 2>&1
ncu -i gpurun_out/hbd_full.ncu-rep --page raw --csv > gpurun_out/hbd_raw.csv
rm -f gpurun_out/hbd_full.ncu-rep
ls -la gpurun_out
