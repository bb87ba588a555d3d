set -x
time (python -m pytest tests/ -x -q -m gpu > gpurun_out/r2t_tests.log 2>&1); tail -2 gpurun_out/r2t_tests.log
time (python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2t_smoke.log 2>&1); tail -1 gpurun_out/r2t_smoke.log
time (python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > gpurun_out/r2t_ref.json 2> gpurun_out/r2t_ref.err); cut -c1-200 gpurun_out/r2t_ref.json
time (python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r2t_bench.json 2> gpurun_out/r2t_bench.err); cut -c1-200 gpurun_out/r2t_bench.json
