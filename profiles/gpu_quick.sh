LT_TORQUE_TC5=1 python profiles/variants.py run 65536 131072 32768 2>&1 | tee gpurun_out/variants_tc5_dyn2.txt
LT_B200_LIB=build/variants/tc5_dyn1.so timeout 600 python -m pytest tests/test_gpu_step.py -m gpu -q -x -k tc5 2>&1 | tail -n 1
LT_TORQUE_TC5=1 LT_B200_LIB=build/variants/tc5_dyn1_trace.so python profiles/tc5_trace.py 65536 2>&1 | tail -n 10
