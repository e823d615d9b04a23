mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gait_corl.py -q > gpurun_out/pytest_gait.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gait.log
