# scratch script for one-off GPU checks (rewritten freely; the full round is profiles/gpu_round.sh)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu.log
LT_TORQUE_TC5=0 timeout 900 python -m pytest tests/test_gpu_step.py -m gpu -x -q > gpurun_out/pytest_gpu_mma.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu_mma.log
LT_TORQUE_TC5=1 timeout 900 python -m pytest tests/test_gpu_step.py tests/test_gpu_env.py -m gpu -x -q > gpurun_out/pytest_gpu_tc5.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu_tc5.log
timeout 300 python bench.py --no-ppo --no-cpu --steps 1000 > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err
