# usage: bash scripts/gpu_profile_r2.sh <tag>   — ncu evidence of the shipped kernels on the default workload (C2, 512 streams):
#   1. launch list of the device-resident pass ALONE (no e2e lanes, no latency probe): per-launch gpu__time_duration
#   2. one `ncu --set full` capture per hot kernel (steady-state launch), with source
TAG=$1
mkdir -p gpurun_out
CMD="python bench.py --steps 4 --warmup 3 --device-only --repeats 1"
$CMD > gpurun_out/${TAG}_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/${TAG}_plain.log; exit 1; }
KR='regex:clahe|pyr_|lk_|shi_tomasi|post_lk|set_mask|select_finish|copy_image|ransac|st_'
ncu --metrics gpu__time_duration.sum --clock-control none -k "$KR" -c 400 --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu_l.log 2>&1; echo "launch list rc=$?"
for K in lk_fast21_kernel shi_tomasi_kernel clahe_lut_kernel clahe_apply_kernel pyr_down_kernel pyr_pad_x_kernel pyr_pad_y_kernel post_lk_ransac_set_mask_kernel select_finish_kernel; do
  SK=4; [ $K = pyr_down_kernel ] && SK=12
  ncu --set full --clock-control none --import-source on -k "regex:^$K" -s $SK -c 1 -o gpurun_out/${TAG}_$K -f $CMD > gpurun_out/${TAG}_ncu_$K.log 2>&1; echo "$K rc=$?"
done
