set -x
timeout 900 python profiles/experiments/sanitize_driver.py --checked > gpurun_out/r2s_checked.log 2>&1; echo "exit=$?" >> gpurun_out/r2s_checked.log
grep -c "QNN_VM_CHECKS: violation" gpurun_out/r2s_checked.log; tail -25 gpurun_out/r2s_checked.log
