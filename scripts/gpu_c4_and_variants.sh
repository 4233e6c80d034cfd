TAG=$1; shift
mkdir -p gpurun_out
python bench.py --config C4 --streams 128 --no-cpu > gpurun_out/${TAG}_bench_c4.json 2> gpurun_out/${TAG}_bench_c4.err; echo "bench c4 rc=$?"; tail -2 gpurun_out/${TAG}_bench_c4.err
bash scripts/gpu_variants.sh $TAG "nothing_selected_xyz" "$@"
