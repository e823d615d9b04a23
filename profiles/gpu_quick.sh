mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_terrain.py -m gpu -x -q > gpurun_out/t_terrain.log 2>&1; echo "rc=$?" >> gpurun_out/t_terrain.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu.log
timeout 600 python bench.py --no-ppo --no-cpu --steps 500 > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err
