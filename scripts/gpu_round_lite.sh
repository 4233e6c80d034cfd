# usage: bash scripts/gpu_round_lite.sh <tag>  — the evidence of scripts/gpu_round_r2.sh without the per-kernel ncu captures that a
# change to ONE kernel does not touch: GPU test log, bench lines (C2 with CPU baseline + C1 sub-record, C1, C3, C4, reference
# arm), launch list, ncu --set full of shi_tomasi and lk only, latency probe
TAG=$1
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/${TAG}_pytest_gpu.txt 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/${TAG}_pytest_gpu.txt
python bench.py > gpurun_out/${TAG}_bench_c2.json 2> gpurun_out/${TAG}_bench_c2.err; echo "bench c2 rc=$?"
python bench.py --config C1 --no-cpu > gpurun_out/${TAG}_bench_c1.json 2> gpurun_out/${TAG}_bench_c1.err; echo "bench c1 rc=$?"
python bench.py --config C3 --streams 256 --no-cpu > gpurun_out/${TAG}_bench_c3.json 2> gpurun_out/${TAG}_bench_c3.err; echo "bench c3 rc=$?"
python bench.py --config C4 --streams 128 --no-cpu > gpurun_out/${TAG}_bench_c4.json 2> gpurun_out/${TAG}_bench_c4.err; echo "bench c4 rc=$?"
python bench.py --impl reference --steps 5 --warmup 2 > gpurun_out/${TAG}_bench_ref.json 2> gpurun_out/${TAG}_bench_ref.err; echo "ref rc=$?"
CMD="python bench.py --steps 4 --warmup 3 --device-only --repeats 1"
KR='regex:clahe|pyr_|lk_|shi_tomasi|post_lk|set_mask|select_finish|copy_image|ransac|st_'
$CMD > gpurun_out/${TAG}_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -k "$KR" -c 400 --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu_l.log 2>&1; echo "launch list rc=$?"
for K in shi_tomasi_kernel lk_fast21_kernel; do
  ncu --set full --clock-control none --import-source on -k "regex:^$K" -s 4 -c 1 -o gpurun_out/${TAG}_$K -f $CMD > gpurun_out/${TAG}_ncu_$K.log 2>&1; echo "$K rc=$?"
done
python scripts/latency_probe.py C2 300 > gpurun_out/${TAG}_latency.txt 2>&1; python scripts/latency_probe.py C1 300 >> gpurun_out/${TAG}_latency.txt 2>&1; tail -4 gpurun_out/${TAG}_latency.txt
