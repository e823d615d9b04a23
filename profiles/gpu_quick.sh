# scratch: which actuator-net kernel is faster at small shard sizes (graph-free eager loop; host-bound below ~14 us per launch)
for t in 0 1; do
for n in 4096 8192 16384 32768; do
echo "TC5=$t $(LT_TORQUE_TC5=$t LT_B200_LIB=build/variants/tc5_base.so python profiles/variants.py one $n)"
done; done
for t in 0 1; do
LT_TORQUE_TC5=$t timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:lt_torque -c 12 --csv --log-file gpurun_out/tq_$t.csv python profiles/run_step.py 8192 6 > /dev/null 2>&1
echo "ncu TC5=$t 8192 envs: $(tail -n 4 gpurun_out/tq_$t.csv | awk -F'","' '{print $NF}' | tr -d '"' | tr '\n' ' ')"
done
