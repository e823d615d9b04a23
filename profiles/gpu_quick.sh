mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_vel_env.py -m gpu -q -x 2>&1 | tail -n 2
python - <<'PY'
import json, bench
for n in (4096, 65536):
    r = bench.cfg4_step_timing("cuda:0", n, steps=40)
    print(n, round(r["ms_per_step"],3), round(r["device_mode"]["ms_per_step"],3), round(r["device_mode"]["graph_replay"]["ms_per_step"],3), round(r["device_mode"]["graph_replay"]["value"]/1e6,1))
PY
