mkdir -p gpurun_out
python - <<'PY' > gpurun_out/cfg4.log 2>&1
import json, torch, bench
for n in (4096, 65536):
    print(json.dumps(bench.cfg4_step_timing("cuda:0", n, steps=40)))
PY
tail -n 12 gpurun_out/cfg4.log
