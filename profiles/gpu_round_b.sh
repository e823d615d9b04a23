# Reduced GPU round (kernels of the cfg 5 step and the PPO path unchanged since profiles/r02_ncu_full.md / r02_gemm_full.md):
# parity tests, smoke, bench (both arms), launch lists of the cfg 5 and cfg 4 steps, --set full of one cfg 4 step.
set -x
TAG=${1:-r02b}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm --format=csv > gpurun_out/smi.txt
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?" >> gpurun_out/bench.err
timeout 300 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
timeout 300 python profiles/run_step.py 65536 6 > gpurun_out/run_step.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches.csv python profiles/run_step.py 65536 6 > gpurun_out/ncu_launch.log 2>&1
for n in 4096 65536; do
timeout 300 python profiles/run_cfg4.py $n 4 > gpurun_out/run_cfg4_$n.log 2>&1 && \
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:lt_ -c 200 --csv --log-file gpurun_out/${TAG}_cfg4_launches_$n.csv python profiles/run_cfg4.py $n 4 > gpurun_out/ncu_cfg4_$n.log 2>&1
done
timeout 900 ncu --set full --clock-control none -k regex:lt_ --launch-skip 50 --launch-count 16 -o /tmp/${TAG}_cfg4_full -f python profiles/run_cfg4.py 65536 4 > gpurun_out/ncu_cfg4_full.log 2>&1
ncu -i /tmp/${TAG}_cfg4_full.ncu-rep --page raw --csv > gpurun_out/${TAG}_cfg4_full_raw.csv 2>/dev/null
du -sh gpurun_out
