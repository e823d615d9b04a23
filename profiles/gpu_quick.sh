mkdir -p gpurun_out
timeout 900 ncu --set full --import-source on --clock-control none -k regex:lt_gemm_kernel -s 156 -c 1 -f -o gpurun_out/gemm_fwd python profiles/run_ppo.py --acts 1 --epochs 1 > gpurun_out/ncu_gemm.log 2>&1
