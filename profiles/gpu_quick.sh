# scratch script for one-off GPU checks (rewritten freely; the full round is profiles/gpu_round.sh)
mkdir -p gpurun_out
LT_TORQUE_TC5=1 timeout 200 python -m pytest tests/test_gpu_step.py -m gpu -x -q -s -k "torque_kernel" > gpurun_out/t_tc5_1.log 2>&1; echo "rc=$?" >> gpurun_out/t_tc5_1.log
: > gpurun_out/q.log
for n in 65536; do
  LT_TORQUE_TC5=1 timeout 300 python bench.py --no-ppo --no-cpu --steps 1000 --envs-per-gpu $n 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('tc5 narrow',$n,round(d['value']/1e6,1),'M/s step_us',round(d['ms_per_step']*1e3,1),'post_us',round(d['post_physics_only']['ms_per_step']*1e3,1))" >> gpurun_out/q.log
done
