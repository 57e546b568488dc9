#!/bin/sh
# Round profile pass on the GPU box (run under gpurun from the repo root): plain bench first,
# then the ncu launch list of the same command, then --set full captures of the kernels.
set -x
python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
  --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 1 --no-cpu --e2e-steps 0 \
  > gpurun_out/ncu_launches.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:txq --launch-skip 19 -c 19 \
  -o gpurun_out/txq_full python bench.py --steps 1 --warmup 1 --no-cpu --no-me --no-extras --e2e-steps 0 \
  > gpurun_out/ncu_txq.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:pel_kernel --launch-skip 2 -c 2 \
  -o gpurun_out/me_full python bench.py --sizes 0 --steps 2 --warmup 1 --no-cpu --no-extras --e2e-steps 0 \
  > gpurun_out/ncu_me.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:hadamard_kernel --launch-skip 6 -c 6 \
  -o gpurun_out/had_full python bench.py --sizes 0 --steps 2 --warmup 1 --no-cpu --no-me --e2e-steps 0 \
  > gpurun_out/ncu_had.log 2>&1
ncu -i gpurun_out/had_full.ncu-rep --page raw --csv > gpurun_out/had_raw.csv
rm -f gpurun_out/had_full.ncu-rep
# the reports are too big to travel back (64 MiB cap): export what profiles/ needs here
ncu -i gpurun_out/txq_full.ncu-rep --page raw --csv > gpurun_out/txq_raw.csv
ncu -i gpurun_out/me_full.ncu-rep --page raw --csv > gpurun_out/me_raw.csv
ncu -i gpurun_out/me_full.ncu-rep --page source --csv --print-source sass > gpurun_out/me_src.csv
ncu -i gpurun_out/txq_full.ncu-rep --page source --csv --print-source sass --kernel-name regex:'txq_kernel<32, 32' > gpurun_out/txq32_src.csv
ncu -i gpurun_out/txq_full.ncu-rep --page source --csv --print-source sass --kernel-name regex:'txq_kernel<64, 64' > gpurun_out/txq64_src.csv
rm -f gpurun_out/txq_full.ncu-rep gpurun_out/me_full.ncu-rep
ls -la gpurun_out
