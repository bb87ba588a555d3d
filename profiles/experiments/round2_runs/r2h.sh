set -x
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/r2h_tests.log 2>&1; echo "tests exit=$?" >> gpurun_out/r2h_tests.log
tail -4 gpurun_out/r2h_tests.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/r2h_bench_cfg3.json 2> gpurun_out/r2h_bench_cfg3.err; echo bench=$?
