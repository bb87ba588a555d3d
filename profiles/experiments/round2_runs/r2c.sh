set -x
python -m pytest tests -m gpu -x -q > gpurun_out/r2c_tests.log 2>&1; echo "tests exit=$?" >> gpurun_out/r2c_tests.log
tail -4 gpurun_out/r2c_tests.log
python -m pytest tests/test_ddp_gpu.py -m gpu -x -q -s 2>&1 | grep -E "DDP-OK|passed|failed|Error" > gpurun_out/r2c_ddp.log; cat gpurun_out/r2c_ddp.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 profiles/cfg5_sweep.py --data-len 2000 --epochs 2 --hdce-epochs 1 --n-per-pair 200 --out gpurun_out/r2c_sweep_small.json > gpurun_out/r2c_sweep_small.log 2>&1; echo sweep=$?
grep -E "^P=|Error|error" gpurun_out/r2c_sweep_small.log | head
