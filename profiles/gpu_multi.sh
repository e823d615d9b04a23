# multi-GPU bench on one box:  bash profiles/gpu_multi.sh N   -- the driver's command line; the JSON line carries the weak figure
# (65,536 envs per GPU) and the strong_65536_total block (65,536 envs in total, PPO iteration with the NCCL exchange timed apart)
mkdir -p gpurun_out
N=${1:-2}
NCCL_DEBUG=INFO timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_${N}gpu.json 2> gpurun_out/bench_${N}gpu.err; echo "rc=$?" >> gpurun_out/bench_${N}gpu.err
grep -c "NCCL INFO" gpurun_out/bench_${N}gpu.err > gpurun_out/bench_${N}gpu.nccl_lines; grep -m2 "nranks\|NVLS" gpurun_out/bench_${N}gpu.err >> gpurun_out/bench_${N}gpu.nccl_lines
tail -c 3000 gpurun_out/bench_${N}gpu.err > gpurun_out/bench_${N}gpu.err.tail; rm gpurun_out/bench_${N}gpu.err
