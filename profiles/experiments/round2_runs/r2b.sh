set -x
python -m pytest tests -m gpu -x -q > gpurun_out/r2b_tests.log 2>&1; echo "tests exit=$?" >> gpurun_out/r2b_tests.log
tail -5 gpurun_out/r2b_tests.log
python -m pytest tests/test_ddp_gpu.py -m gpu -x -q -s > gpurun_out/r2b_ddp.log 2>&1; echo "ddp exit=$?" >> gpurun_out/r2b_ddp.log
grep -E "DDP-OK|passed|failed|exit" gpurun_out/r2b_ddp.log
for n in 13 14 16; do python profiles/experiments/prof_n.py $n 3; done 2>&1 | tee gpurun_out/r2b_prof_n.log
python bench.py --steps 10 --warmup 3 > gpurun_out/r2b_bench_cfg3.json 2> gpurun_out/r2b_bench_cfg3.err; echo bench1=$?
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r2b_bench_cfg3_2gpu.json 2> gpurun_out/r2b_bench_cfg3_2gpu.err; echo bench2=$?
tail -3 gpurun_out/r2b_bench_cfg3.err gpurun_out/r2b_bench_cfg3_2gpu.err
