set -x
timeout 300 python -m pytest tests/test_qnn_gpu.py -m gpu -x -q -k "largest or config4 or deep_cluster or cta_per_sample or multi_warp" > gpurun_out/r2f_tests.log 2>&1; echo "tests exit=$?" >> gpurun_out/r2f_tests.log
tail -6 gpurun_out/r2f_tests.log
for n in 14 15 16; do timeout 120 python profiles/experiments/prof_n.py $n 3; done 2>&1 | tee gpurun_out/r2f_prof_n.log
timeout 120 python profiles/experiments/prof_n.py 16 8 576 2>&1 | tee -a gpurun_out/r2f_prof_n.log
timeout 300 python profiles/experiments/sanitize_driver.py > gpurun_out/r2f_driver.log 2>&1; tail -3 gpurun_out/r2f_driver.log
