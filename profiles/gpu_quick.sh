mkdir -p gpurun_out
python bench.py --steps 500 --warmup 5 --no-ppo --no-cpu --e2e-chunks 1 2>gpurun_out/bench_err.log > gpurun_out/bench_quick.json
python -c "
import json
d=json.loads(open('gpurun_out/bench_quick.json').read().strip().splitlines()[-1]); print(d['value'], d['ms_per_step']); print(json.dumps(d['roofline'])[:900])"
