# One full GPU round on a B200 box: parity tests, smoke, bench (both arms), ncu launch lists and --set full captures.
# Usage (from the repo root on the box):  bash profiles/gpu_round.sh [tag]       outputs -> gpurun_out/
set -x
TAG=${1:-r02}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm --format=csv > gpurun_out/smi.txt
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?" >> gpurun_out/bench.err
timeout 300 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
# launch lists (the programs have just exited 0 without ncu): the env step, then the PPO path
timeout 300 python profiles/run_step.py 65536 6 > gpurun_out/run_step.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches.csv python profiles/run_step.py 65536 6 > gpurun_out/ncu_launch.log 2>&1
timeout 300 python profiles/run_ppo.py --acts 1 > gpurun_out/run_ppo.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${TAG}_ppo_launches.csv python profiles/run_ppo.py --acts 1 > gpurun_out/ncu_ppo_launch.log 2>&1
# --set full: one full env step (4 torque + step + compact + obs), and the 39 GEMMs of one steady-state minibatch (no source: size)
timeout 900 ncu --set full --clock-control none --import-source on -k regex:lt_ --launch-skip 14 --launch-count 7 -o gpurun_out/${TAG}_full -f python profiles/run_step.py 65536 4 > gpurun_out/ncu_full.log 2>&1
ncu -i gpurun_out/${TAG}_full.ncu-rep --page raw --csv > gpurun_out/${TAG}_full_raw.csv 2>/dev/null
timeout 900 ncu --set full --clock-control none -k regex:lt_gemm_kernel -s 153 -c 39 -o /tmp/${TAG}_gemm_full -f python profiles/run_ppo.py --acts 1 --epochs 1 > gpurun_out/ncu_gemm_full.log 2>&1
ncu -i /tmp/${TAG}_gemm_full.ncu-rep --page raw --csv > gpurun_out/${TAG}_gemm_full_raw.csv 2>/dev/null      # the 39-kernel report itself is > 64 MiB
du -sh gpurun_out
ls -la gpurun_out
