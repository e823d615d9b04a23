mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_terrain.py -m gpu -x -q > gpurun_out/t_terrain.log 2>&1; echo "rc=$?" >> gpurun_out/t_terrain.log
python - > gpurun_out/terrain_time.log 2>&1 <<'PY'
import torch, bench
print(bench.terrain_build_timing("cuda:0"))
PY
