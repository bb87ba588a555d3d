set -x
mkdir -p /tmp/rep
CMD="python bench.py --steps 3 --warmup 3 --no-cpu --no-secondary"
$CMD > gpurun_out/r2q_plain.json 2> gpurun_out/r2q_plain.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2q_launches.csv $CMD > gpurun_out/r2q_ncu_launches.log 2>&1
cap() {  # name, args...
  name=$1; shift
  python profiles/experiments/prof_n.py "$@" > gpurun_out/r2q_p_$name.log 2>&1 && \
  ncu --set full --clock-control none -k 'regex:^k_vm$' --launch-skip 2 --launch-count 2 -o /tmp/rep/$name python profiles/experiments/prof_n.py "$@" > gpurun_out/r2q_ncu_$name.log 2>&1
  python profiles/summarize.py full /tmp/rep/$name.ncu-rep > gpurun_out/r2q_full_$name.txt 2>&1
}
cap cfg3 10 3 65536
cap cfg4 16 8 576
cap n13 13 3
cap n14 14 3
cap cfg2 6 3 4096 narrow
cat gpurun_out/r2q_p_*.log
