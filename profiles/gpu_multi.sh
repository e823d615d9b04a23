# multi-GPU bench on one box:  bash profiles/gpu_multi.sh N   (weak scaling at 65,536 envs/GPU, then 65,536 envs in total)
mkdir -p gpurun_out
N=${1:-2}
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 1000 --warmup 10 --no-cpu > gpurun_out/bench_${N}gpu.json 2> gpurun_out/bench_${N}gpu.err; echo "rc=$?" >> gpurun_out/bench_${N}gpu.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus $N --steps 2000 --warmup 10 --no-cpu --no-ppo --envs-per-gpu $((65536 / N)) > gpurun_out/bench_${N}gpu_65536total.json 2> gpurun_out/bench_${N}gpu_65536total.err; echo "rc=$?" >> gpurun_out/bench_${N}gpu_65536total.err
