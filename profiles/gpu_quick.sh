# scratch script for one-off GPU checks (rewritten freely; the full round is profiles/gpu_round.sh)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
timeout 600 python bench.py --steps 50 --warmup 3 > gpurun_out/bench_short.json 2> gpurun_out/bench_short.err; echo "rc=$?" >> gpurun_out/bench_short.err
