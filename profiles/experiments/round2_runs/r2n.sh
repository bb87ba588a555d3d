set -x
timeout 600 python -m pytest tests/test_module_gpu.py -m gpu -x -q > gpurun_out/r2n_tests.log 2>&1; tail -3 gpurun_out/r2n_tests.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/r2n_bench_cfg3.json 2> gpurun_out/r2n_bench_cfg3.err; echo bench=$?; tail -3 gpurun_out/r2n_bench_cfg3.err
