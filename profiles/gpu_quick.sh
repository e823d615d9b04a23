mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_pipeline.py -m gpu -x -q > gpurun_out/t_pipe.log 2>&1; echo "rc=$?" >> gpurun_out/t_pipe.log

: > gpurun_out/q.log
for c in 2 4 8; do
  timeout 300 python bench.py --no-ppo --no-cpu --steps 300 --e2e-chunks $c 2>gpurun_out/bench_e2e_$c.err | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('chunks',$c,'e2e',round(d['e2e']['value']/1e6,2),'ms',round(d['e2e']['ms_per_step'],3),'single',round(d['e2e']['single_stream']['value']/1e6,2))" >> gpurun_out/q.log
done
