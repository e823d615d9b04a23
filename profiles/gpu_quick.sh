python profiles/variants.py run 65536 16384 2>&1 | tee gpurun_out/variants_step.txt
