#!/bin/sh
# Round-2 closing evidence (run under gpurun from the repo root): GPU tests, smoke(), both bench arms, the ncu
# launch list of the bench command and one --set full capture of the full-pel motion search with its
# per-region stall summary.  Every ncu pass runs only after the same command exited 0 without ncu.
set -x
python -m pytest tests -x -q -m gpu > gpurun_out/r2_pytest_gpu_final.log 2>&1; tail -2 gpurun_out/r2_pytest_gpu_final.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r2_smoke.log 2>&1; tail -1 gpurun_out/r2_smoke.log
python bench.py --impl reference > gpurun_out/r2_bench_reference_n1_final.json 2> gpurun_out/ref.err
python bench.py > gpurun_out/r2_bench_n1_final.json 2> gpurun_out/bench.err || exit 1
python bench.py --steps 2 --warmup 1 --no-cpu --e2e-steps 0 > gpurun_out/plain.json 2> gpurun_out/plain.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2_v3_launches.csv \
  python bench.py --steps 2 --warmup 1 --no-cpu --e2e-steps 0 > gpurun_out/ncu_launches.log 2>&1
sh tools/profile_me_src.sh > gpurun_out/profile_me_src.log 2>&1
python tools/ncu_table.py gpurun_out/me_src_raw.csv > gpurun_out/r2_me_v5_ncu_table.md 2>&1
ls -la gpurun_out
