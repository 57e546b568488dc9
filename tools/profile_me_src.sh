#!/bin/sh
# source-level ncu pass over the full-pel kernel on the workload of tools/me_phase_timing.py ("$@" = its options:
# none = 4 frames of configs[2], --hbd = the 10-bit searches of configs[4], --tf = the temporal filter's);
# summarised on the box (the CSV is too large to travel).  TAG names the output files.
TAG=${TAG:-me}
python tools/me_phase_timing.py "$@" > gpurun_out/${TAG}_plain.txt 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:fullpel_kernel --launch-skip 1 -c 1 \
  -o /tmp/${TAG}_src python tools/me_phase_timing.py "$@" > gpurun_out/ncu_${TAG}_src.log 2>&1
ncu -i /tmp/${TAG}_src.ncu-rep --page source --csv --print-source sass > /tmp/${TAG}_src.csv 2> gpurun_out/${TAG}_src.err
python tools/src_regions.py /tmp/${TAG}_src.csv fullpel 100 60 > gpurun_out/${TAG}_src_regions.txt 2>&1
ncu -i /tmp/${TAG}_src.ncu-rep --page raw --csv > gpurun_out/${TAG}_src_raw.csv 2>/dev/null
head -c 3000 gpurun_out/${TAG}_src_regions.txt
