mkdir -p gpurun_out
set -x
python bench.py > gpurun_out/bench_v0.json 2> gpurun_out/bench_v0.err; echo "rc=$?"; tail -c 3000 gpurun_out/bench_v0.json; tail -5 gpurun_out/bench_v0.err
python bench.py --config C1 --no-cpu > gpurun_out/bench_v0_c1.json 2>> gpurun_out/bench_v0.err; echo "rc=$?"
KR='regex:clahe|pyr_down|lk_kernel|eig_kernel|post_lk|set_mask|fill_mask|paint_mask|select_finish|copy_image'
python bench.py --steps 3 --warmup 3 --no-cpu --latency-frames 5 > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -k "$KR" -c 300 --csv --log-file gpurun_out/launches_v0.csv python bench.py --steps 3 --warmup 3 --no-cpu --latency-frames 5 > gpurun_out/ncu_v0.log 2>&1; echo "ncu rc=$?"; tail -3 gpurun_out/ncu_v0.log
