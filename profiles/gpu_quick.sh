mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_vel_env.py -m gpu -q -x 2>&1 | tail -n 3
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r02_cfg4_launches.csv python profiles/run_cfg4.py 65536 4 > gpurun_out/ncu_cfg4.log 2>&1
grep "lt_vel_observe" gpurun_out/r02_cfg4_launches.csv | tail -n 2
