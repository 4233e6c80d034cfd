# single-stream (latency mode) ncu captures of the glue kernels with the finest PC-sampling interval
for K in ${KERNELS:-set_mask_kernel select_finish_kernel}; do
ncu --set full --warp-sampling-interval 0 --clock-control none --import-source on -k "regex:^$K" -s 30 -c 1 -o gpurun_out/lat2_$K -f python scripts/latency_probe.py ${CFG:-C1} 40 > gpurun_out/ncu_lat2_$K.log 2>&1; tail -1 gpurun_out/ncu_lat2_$K.log
done
