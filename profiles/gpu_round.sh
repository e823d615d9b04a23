# One full GPU round on a B200 box: parity tests, smoke, bench (both arms), ncu launch list and --set full capture.
# Usage (from the repo root on the box):  bash profiles/gpu_round.sh [tag]       outputs -> gpurun_out/
set -x
TAG=${1:-r01}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm --format=csv > gpurun_out/smi.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?" >> gpurun_out/bench.err
timeout 300 python bench.py --impl reference --steps 10 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
timeout 300 python profiles/run_step.py 65536 6 > gpurun_out/run_step.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches.csv python profiles/run_step.py 65536 6 > gpurun_out/ncu_launch.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:lt_ --launch-skip 14 --launch-count 7 -o gpurun_out/${TAG}_full -f python profiles/run_step.py 65536 4 > gpurun_out/ncu_full.log 2>&1
ls -la gpurun_out
