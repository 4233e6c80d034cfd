# usage: bash scripts/gpu_round.sh <tag>   — everything the judge reads for one round state: tests, bench lines (C2 default with CPU
# baseline, C1), reference arm, launch list and ncu --set full of the hot kernels (C2 default workload)
TAG=$1
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -3
python bench.py > gpurun_out/${TAG}_bench_c2.json 2> gpurun_out/${TAG}_bench_c2.err; echo "bench c2 rc=$?"
python bench.py --config C1 --no-cpu > gpurun_out/${TAG}_bench_c1.json 2> gpurun_out/${TAG}_bench_c1.err; echo "bench c1 rc=$?"
python bench.py --config C3 --streams 256 --no-cpu > gpurun_out/${TAG}_bench_c3.json 2> gpurun_out/${TAG}_bench_c3.err; echo "bench c3 rc=$?"
python bench.py --config C4 --streams 128 --no-cpu > gpurun_out/${TAG}_bench_c4.json 2> gpurun_out/${TAG}_bench_c4.err; echo "bench c4 rc=$?"
python bench.py --impl reference --steps 5 --warmup 2 > gpurun_out/${TAG}_bench_ref.json 2> gpurun_out/${TAG}_bench_ref.err; echo "ref rc=$?"
CMD="python bench.py --steps 3 --warmup 3 --no-cpu --latency-frames 1"
KR='regex:clahe|pyr_down|lk_|shi_tomasi|post_lk|set_mask|select_finish|copy_image|ransac'
$CMD > gpurun_out/${TAG}_plain.log 2>&1 || { echo "plain run failed"; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -k "$KR" -c 400 --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu_l.log 2>&1; echo "launch list rc=$?"
for K in lk_fast21_kernel shi_tomasi_kernel clahe_lut_kernel clahe_apply_kernel pyr_down_kernel post_lk_ransac_set_mask_kernel select_finish_kernel; do
  SK=4; [ $K = pyr_down_kernel ] && SK=12
  ncu --set full --clock-control none --import-source on -k "regex:^$K" -s $SK -c 1 -o gpurun_out/${TAG}_$K -f $CMD > gpurun_out/${TAG}_ncu_$K.log 2>&1; echo "$K rc=$?"
done
# latency mode (one stream, programmatic dependent launches): p50 + per-kernel split, with and without the RANSAC
python scripts/latency_probe.py C2 300 > gpurun_out/${TAG}_latency.txt 2>&1; python scripts/latency_probe.py C1 300 >> gpurun_out/${TAG}_latency.txt 2>&1
python scripts/latency_probe.py C2 300 off >> gpurun_out/${TAG}_latency.txt 2>&1; tail -6 gpurun_out/${TAG}_latency.txt
