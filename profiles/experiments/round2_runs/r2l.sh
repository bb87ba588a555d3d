set -x
timeout 600 python -m pytest tests/test_qnn_gpu.py -m gpu -x -q -k "narrow or variant or sixteen" > gpurun_out/r2l_tests.log 2>&1; tail -2 gpurun_out/r2l_tests.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/r2l_bench_cfg3.json 2> gpurun_out/r2l_bench_cfg3.err; echo bench=$?
python profiles/experiments/prof_n.py 10 3 65536 16-per-thread > gpurun_out/r2l_prof_k4.log 2>&1 && \
ncu --set full --clock-control none -k 'regex:^k_vm$' --launch-skip 2 --launch-count 2 -o gpurun_out/r2l_k4_n10 python profiles/experiments/prof_n.py 10 3 65536 16-per-thread > gpurun_out/r2l_ncu_k4.log 2>&1
cat gpurun_out/r2l_prof_k4.log
python profiles/experiments/prof_n.py 10 3 65536 wide > gpurun_out/r2l_prof_k5.log 2>&1 && \
ncu --set full --clock-control none -k 'regex:^k_vm$' --launch-skip 2 --launch-count 2 -o gpurun_out/r2l_k5_n10 python profiles/experiments/prof_n.py 10 3 65536 wide > gpurun_out/r2l_ncu_k5.log 2>&1
cat gpurun_out/r2l_prof_k5.log
