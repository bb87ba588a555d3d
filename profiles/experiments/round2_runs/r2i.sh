set -x
timeout 900 python -m pytest tests/test_qnn_gpu.py -m gpu -x -q > gpurun_out/r2i_tests.log 2>&1; echo "tests exit=$?" >> gpurun_out/r2i_tests.log
tail -5 gpurun_out/r2i_tests.log
timeout 600 python profiles/experiments/variants.py > gpurun_out/r2i_variants.log 2>&1; cat gpurun_out/r2i_variants.log
