mkdir -p gpurun_out
LT_TORQUE_TC5=1 timeout 120 python -m pytest tests/test_gpu_step.py -m gpu -x -q -s -k "torque_kernel" > gpurun_out/t_tc5_1.log 2>&1; echo "rc=$?" >> gpurun_out/t_tc5_1.log
nvidia-smi --query-gpu=name --format=csv,noheader > gpurun_out/alive.txt 2>&1
: > gpurun_out/q.log
for n in 65536; do for v in 1; do
  LT_TORQUE_TC5=$v timeout 300 python bench.py --no-ppo --no-cpu --steps 1000 --envs-per-gpu $n 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('tc5',$v,$n,round(d['value']/1e6,1),'M/s step_us',round(d['ms_per_step']*1e3,1),'post_us',round(d['post_physics_only']['ms_per_step']*1e3,1))" >> gpurun_out/q.log
done; done
