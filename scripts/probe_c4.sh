python scripts/latency_probe.py C4 30 > gpurun_out/probe_c4.log 2>&1; tail -2 gpurun_out/probe_c4.log
ncu --set full --clock-control none --import-source on -k "regex:^set_mask_kernel" -s 20 -c 1 -o gpurun_out/c4_set_mask -f python scripts/latency_probe.py C4 30 > gpurun_out/ncu_c4.log 2>&1; tail -2 gpurun_out/ncu_c4.log
