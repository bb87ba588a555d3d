set -x
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/r2p_tests.log 2>&1; echo "tests exit=$?" >> gpurun_out/r2p_tests.log
tail -3 gpurun_out/r2p_tests.log
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/r2p_bench_cfg3.json 2> gpurun_out/r2p_bench_cfg3.err; echo bench=$?
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29516 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r2p_bench_cfg3_2gpu.json 2> gpurun_out/r2p_bench_cfg3_2gpu.err; echo bench2=$?
for n in 11 12 13 14 15 16; do timeout 120 python profiles/experiments/prof_n.py $n 3; done 2>&1 | grep "^n=" | tee gpurun_out/r2p_prof_n.log
