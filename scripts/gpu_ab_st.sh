# usage: bash scripts/gpu_ab_st.sh <tag>  — Shi-Tomasi A/B: GPU tests on the working tree's library, then C2 / C1 bench lines
# (device leg + latency, no CPU arm) for: the tree's library, the same with VS_ST_SKIP=0 (every row step walked), and the
# previous commit's library (_variants/libvs_base.so, built from HEAD with csrc/build.sh).
TAG=$1
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest_gpu.txt 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/${TAG}_pytest_gpu.txt
run() {  # name, env...
  local name=$1; shift
  for CFG in C2 C1; do
    env "$@" python bench.py --config $CFG --no-cpu --no-c1 --latency-frames 100 > gpurun_out/${TAG}_${name}_${CFG}.json 2> gpurun_out/${TAG}_${name}_${CFG}.err; echo "$name $CFG rc=$?"
    python scripts/bench_brief.py $name-$CFG < gpurun_out/${TAG}_${name}_${CFG}.json 2>&1 | tail -2
  done
}
run new VS_NOP=1
run noskip VS_ST_SKIP=0
[ -f _variants/libvs_base.so ] && run base VS_TRACKER_LIB=$PWD/_variants/libvs_base.so
