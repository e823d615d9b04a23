mkdir -p gpurun_out
timeout 900 python profiles/phase_probe.py 65536 300 > gpurun_out/phase_probe.log 2>&1
