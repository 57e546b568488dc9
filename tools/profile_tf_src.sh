python tools/me_phase_timing.py --tf > gpurun_out/tf_plain.txt 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:fullpel_kernel --launch-skip 1 -c 1 -o /tmp/tf_src python tools/me_phase_timing.py --tf > gpurun_out/ncu_tf_src.log 2>&1
ncu -i /tmp/tf_src.ncu-rep --page source --csv --print-source sass > /tmp/tf_src.csv 2> gpurun_out/tf_src.err
python tools/src_regions.py /tmp/tf_src.csv fullpel 100 50 > gpurun_out/tf_src_regions.txt 2>&1
ncu -i /tmp/tf_src.ncu-rep --page raw --csv > gpurun_out/tf_src_raw.csv 2>/dev/null
python tools/ncu_table.py gpurun_out/tf_src_raw.csv
