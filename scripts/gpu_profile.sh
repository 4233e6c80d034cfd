# usage: bash scripts/gpu_profile.sh <tag> [bench args]  — launch list + ncu --set full of each hot kernel
mkdir -p gpurun_out
TAG=$1; shift
CMD="python bench.py --steps 3 --warmup 3 --no-cpu --latency-frames 1 $@"
KR='regex:clahe|pyr_down|lk_|shi_tomasi|post_lk|set_mask|select_finish|copy_image|ransac'
$CMD > gpurun_out/plain_$TAG.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain_$TAG.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -k "$KR" -c 400 --csv --log-file gpurun_out/launches_$TAG.csv $CMD > gpurun_out/ncu_l_$TAG.log 2>&1; echo "launch list rc=$?"
# steady-state launches: skip the first 4 steps' worth
for K in ${KERNELS:-lk_fast21_kernel shi_tomasi_kernel clahe_lut_kernel clahe_apply_kernel pyr_down_kernel set_mask_kernel select_finish_kernel}; do
  SK=4; [ $K = pyr_down_kernel ] && SK=12
  ncu --set full --clock-control none --import-source on -k "regex:^$K" -s $SK -c 1 -o gpurun_out/${TAG}_$K -f $CMD > gpurun_out/ncu_${TAG}_$K.log 2>&1; echo "$K rc=$?"
done
ls -la gpurun_out/*.ncu-rep
