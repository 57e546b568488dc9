#!/bin/sh
# source-level ncu pass over the full-pel kernel on the 4-frame configs[2] workload of tools/me_phase_timing.py;
# summarised on the box (the CSV is too large to travel)
python tools/me_phase_timing.py > gpurun_out/me_plain.txt 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:fullpel_kernel --launch-skip 1 -c 1 \
  -o /tmp/me_src python tools/me_phase_timing.py > gpurun_out/ncu_me_src.log 2>&1
ncu -i /tmp/me_src.ncu-rep --page source --csv --print-source sass > /tmp/me_src.csv 2> gpurun_out/me_src.err
python tools/src_regions.py /tmp/me_src.csv fullpel 100 60 > gpurun_out/me_src_regions.txt 2>&1
ncu -i /tmp/me_src.ncu-rep --page raw --csv > gpurun_out/me_src_raw.csv 2>/dev/null
head -c 3000 gpurun_out/me_src_regions.txt
