set -x
python profiles/experiments/sanitize_driver.py > gpurun_out/r2e_plain.log 2>&1 && \
timeout 1200 compute-sanitizer --tool memcheck --print-limit 20 python profiles/experiments/sanitize_driver.py > gpurun_out/r2e_memcheck.log 2>&1; echo memcheck=$?
tail -15 gpurun_out/r2e_memcheck.log
