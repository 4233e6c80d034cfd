for K in set_mask_kernel select_finish_kernel; do
ncu --set full --clock-control none --import-source on -k "regex:^$K" -s 30 -c 1 -o gpurun_out/lat2_$K -f python scripts/latency_probe.py C1 40 > gpurun_out/ncu_lat2_$K.log 2>&1; tail -1 gpurun_out/ncu_lat2_$K.log
done
