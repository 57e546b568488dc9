python -m pytest tests/test_me_gpu.py -m gpu -x -q 2>&1 | tail -15 > gpurun_out/me_tests.log
python bench.py --sizes 0 --no-cpu --e2e-steps 0 > gpurun_out/bench_me.json 2> gpurun_out/bench_me.err
B200_ME_DEBUG=1 B200AV1_LIB=$PWD/aom-av1-lavish_b200/libb200av1_timing.so python bench.py --sizes 0 --no-cpu --e2e-steps 0 > gpurun_out/bench_me_t.json 2> gpurun_out/bench_me_t.err
