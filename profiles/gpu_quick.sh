mkdir -p gpurun_out
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_8192.csv python profiles/run_step.py 8192 6 > gpurun_out/ncu_launch.log 2>&1
