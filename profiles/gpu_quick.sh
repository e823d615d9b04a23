mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_planner.py tests/test_abi.py -q > gpurun_out/pytest_planner.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_planner.log
