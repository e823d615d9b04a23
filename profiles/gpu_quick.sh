mkdir -p gpurun_out
python - <<'PY'
import json, bench
print(json.dumps(bench.cfg4_ppo_iteration("cuda:0", 4096)))
PY
