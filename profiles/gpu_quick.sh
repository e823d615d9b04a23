timeout 900 python -m pytest tests/test_vel_env.py tests/test_gait_corl.py tests/test_commands.py -m gpu -q -x 2>&1 | tail -n 4
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -n 2
python - <<PY
import bench
for n in (4096, 65536):
    r = bench.cfg4_step_timing("cuda:0", n, steps=40)
    print(n, round(r["ms_per_step"],3), round(r["device_mode"]["ms_per_step"],3), round(r["device_mode"]["graph_replay"]["ms_per_step"],4), round(r["device_mode"]["graph_replay"]["value"]/1e6,1))
PY
