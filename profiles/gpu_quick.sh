mkdir -p gpurun_out
rm -f gpurun_out/chunks.log
for c in 4096,12288,16384,16384,16384 8192,8192,8192,8192,8192,8192,8192,8192 4096,4096,8192,8192,8192,8192,8192,8192,8192 2048,6144,8192,8192,8192,8192,8192,8192,8192; do
  python bench.py --steps 300 --warmup 5 --no-ppo --no-cpu --e2e-shards $c 2>gpurun_out/bench_err.log | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('shards $c', d['e2e']['value'], d['e2e']['ms_per_step'], d['value'])" >> gpurun_out/chunks.log 2>&1
done
tail -n 5 gpurun_out/bench_err.log
