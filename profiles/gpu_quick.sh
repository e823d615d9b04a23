for r in 1 2 4 8 12 24; do
LT_CORL_ROWS=$r timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:lt_corl -c 4 --csv --log-file gpurun_out/corl_$r.csv python profiles/run_cfg4.py 65536 4 > gpurun_out/corl_$r.log 2>&1
echo "rows $r: $(tail -n 2 gpurun_out/corl_$r.csv | awk -F'","' '{print $NF}' | tr -d '"' | tr '\n' ' ')"
done
