python profiles/variants.py run 65536 2>&1 | tee gpurun_out/variants_step3.txt
