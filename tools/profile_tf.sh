#!/bin/sh
# ncu --set full capture of the four chained motion-search launches of one temporal-filter frame pair
set -x
B="python bench.py --only tf"
ncu --set full --clock-control none -k regex:pel_kernel --launch-skip 4 -c 4 -o gpurun_out/tf_full $B > gpurun_out/ncu_tf.log 2>&1
ncu -i gpurun_out/tf_full.ncu-rep --page raw --csv > gpurun_out/tf_raw.csv
rm -f gpurun_out/tf_full.ncu-rep
