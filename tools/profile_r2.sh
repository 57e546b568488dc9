#!/bin/sh
# Round-2 ncu pass (run under gpurun from the repo root, after bench.py has exited 0 without ncu):
# --set full captures of the pixel-input transform kernels (config[3]), the inverse-transform kernels
# and the motion-search kernels (8-bit and the 4K 10-bit ones of configs[4]); the reports are exported
# to CSV here because they are too big to travel.
set -x
B="python bench.py --sizes 0 --steps 2 --warmup 1 --no-cpu --e2e-steps 0 --config5-frames 4"
$B > gpurun_out/prof_plain.json 2> gpurun_out/prof_plain.err || exit 1
ncu --set full --clock-control none --import-source on -k 'regex:txq.*, 2>' --launch-skip 19 -c 19 \
  -o gpurun_out/hbd_full $B --no-me > gpurun_out/ncu_hbd.log 2>&1
ncu -i gpurun_out/hbd_full.ncu-rep --page raw --csv > gpurun_out/hbd_raw.csv
rm -f gpurun_out/hbd_full.ncu-rep
ncu --set full --clock-control none --import-source on -k regex:itx_kernel --launch-skip 3 -c 73 \
  -o gpurun_out/itx_full $B --no-me > gpurun_out/ncu_itx.log 2>&1
ncu -i gpurun_out/itx_full.ncu-rep --page raw --csv > gpurun_out/itx_raw.csv
rm -f gpurun_out/itx_full.ncu-rep
ncu --set full --clock-control none --import-source on -k regex:pel_kernel -c 14 \
  -o gpurun_out/me_full $B > gpurun_out/ncu_me.log 2>&1
ncu -i gpurun_out/me_full.ncu-rep --page raw --csv > gpurun_out/me_raw.csv
# per-SASS-line counters of the first 8-bit full-pel / sub-pel pair only, summarised here (the CSV is ~100 MB)
ncu -i gpurun_out/me_full.ncu-rep --page source --csv --print-source sass > /tmp/me_src.csv 2> gpurun_out/me_src.err
python tools/sass_hot.py /tmp/me_src.csv 30 > gpurun_out/me_sass_hot.txt 2>&1
rm -f gpurun_out/me_full.ncu-rep /tmp/me_src.csv
ls -la gpurun_out
