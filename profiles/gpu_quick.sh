mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_policy.py tests/test_gpu_ppo.py tests/test_zz_no_alloc.py -m gpu -q -x > gpurun_out/pytest_policy.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_policy.log
timeout 300 python profiles/run_ppo.py --reps 2 > gpurun_out/run_ppo.log 2>&1
