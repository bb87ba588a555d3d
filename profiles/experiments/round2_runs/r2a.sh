set -x
python -m pytest tests -m gpu -x -q > gpurun_out/r2a_tests.log 2>&1; echo "tests exit=$?" >> gpurun_out/r2a_tests.log
tail -5 gpurun_out/r2a_tests.log
for n in 13 14 16; do
  SRC=""; if [ $n = 16 ]; then SRC="--import-source on"; fi
  python profiles/experiments/prof_n.py $n 3 > gpurun_out/r2a_prof_n$n.log 2>&1 && \
  ncu --set full --clock-control none $SRC -k 'regex:^k_vm$' --launch-skip 2 --launch-count 2 -o gpurun_out/r2a_n$n python profiles/experiments/prof_n.py $n 3 > gpurun_out/r2a_ncu_n$n.log 2>&1
  cat gpurun_out/r2a_prof_n$n.log; tail -2 gpurun_out/r2a_ncu_n$n.log
done
python bench.py --steps 10 --warmup 3 > gpurun_out/r2a_bench_cfg3.json 2> gpurun_out/r2a_bench_cfg3.err; tail -c 600 gpurun_out/r2a_bench_cfg3.json
