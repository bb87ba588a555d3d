set -x
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/r2m_tests.log 2>&1; echo "tests exit=$?" >> gpurun_out/r2m_tests.log
tail -3 gpurun_out/r2m_tests.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/r2m_bench_cfg3.json 2> gpurun_out/r2m_bench_cfg3.err; echo bench=$?
