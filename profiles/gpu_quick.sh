mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_planner.py -q -x > gpurun_out/pytest_planner.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_planner.log
