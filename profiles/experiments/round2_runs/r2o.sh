set -x
timeout 600 python -m pytest tests/test_qnn_gpu.py tests/test_random_circuits.py -m gpu -x -q -k "largest or config4 or deep_cluster or random_circuits_on_the_gpu" > gpurun_out/r2o_tests.log 2>&1; tail -3 gpurun_out/r2o_tests.log
for n in 14 15 16; do timeout 120 python profiles/experiments/prof_n.py $n 3; done
timeout 120 python profiles/experiments/prof_n.py 16 8 576
