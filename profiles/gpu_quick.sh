mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu.log
: > gpurun_out/q.log
for n in 8192 65536; do
  timeout 300 python bench.py --no-ppo --no-cpu --steps 1000 --envs-per-gpu $n 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print($n,round(d['value']/1e6,1),'M/s step_us',round(d['ms_per_step']*1e3,1),'post_us',round(d['post_physics_only']['ms_per_step']*1e3,1),'e2e',round(d['e2e']['value']/1e6,2))" >> gpurun_out/q.log
done
timeout 300 python bench.py --no-ppo --no-cpu --steps 1000 --envs-per-gpu 8192 --epw 4 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('epw4 8192',round(d['value']/1e6,1),'M/s step_us',round(d['ms_per_step']*1e3,1),'post_us',round(d['post_physics_only']['ms_per_step']*1e3,1))" >> gpurun_out/q.log
