LT_TORQUE_TC5=1 python profiles/variants.py run 65536 131072 2>&1 | tee gpurun_out/variants_tc5_d.txt
