set -x
python -m pytest tests -m gpu -q > gpurun_out/r2d_tests.log 2>&1; echo "tests exit=$?" >> gpurun_out/r2d_tests.log
tail -4 gpurun_out/r2d_tests.log
python -m pytest tests/test_ddp_gpu.py -m gpu -x -q -s 2>&1 | grep -E "DDP-OK|passed|failed|Error" > gpurun_out/r2d_ddp.log; cat gpurun_out/r2d_ddp.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29513 profiles/cfg5_sweep.py --out gpurun_out/r02_cfg5_sweep.json > gpurun_out/r2d_sweep.log 2>&1; echo sweep=$?
grep -E "^P=|Error|error|wrote" gpurun_out/r2d_sweep.log | head
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29514 bench.py --gpus 8 --steps 10 --warmup 3 > gpurun_out/r2d_bench_cfg3_8gpu.json 2> gpurun_out/r2d_bench_cfg3_8gpu.err; echo bench8=$?
