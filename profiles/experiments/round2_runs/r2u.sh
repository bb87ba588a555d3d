set -x
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/r2u_tests.log 2>&1; echo "tests exit=$?" >> gpurun_out/r2u_tests.log
tail -3 gpurun_out/r2u_tests.log
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/r2u_bench_cfg3.json 2> gpurun_out/r2u_bench_cfg3.err; echo bench=$?
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29518 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r2u_bench_cfg3_2gpu.json 2> gpurun_out/r2u_bench_cfg3_2gpu.err; echo bench2=$?
