set -x
python profiles/experiments/prof_n.py 16 3 > gpurun_out/r2g_prof_n16.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k 'regex:^k_vm$' --launch-skip 2 --launch-count 2 -o gpurun_out/r2g_n16 python profiles/experiments/prof_n.py 16 3 > gpurun_out/r2g_ncu_n16.log 2>&1
cat gpurun_out/r2g_prof_n16.log; tail -2 gpurun_out/r2g_ncu_n16.log
