# usage: bash scripts/gpu_ab_quick.sh <tag> [pytest -k expression]  — quick A/B of the tree's library against _variants/libvs_base.so:
# the Shi-Tomasi related GPU tests, C2 / C1 device-leg bench lines of both, one ncu --set full capture of the tree's shi_tomasi
TAG=$1; KEXPR=${2:-"good_features or min_eig or sequence or golden or two_tier or edge or ref_pin"}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "$KEXPR" > gpurun_out/${TAG}_pytest_gpu.txt 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/${TAG}_pytest_gpu.txt
run() {  # name, env...
  local name=$1; shift
  for CFG in C2 C1; do
    env "$@" python bench.py --config $CFG --no-cpu --no-c1 --device-only > gpurun_out/${TAG}_${name}_${CFG}.json 2> gpurun_out/${TAG}_${name}_${CFG}.err; echo "$name $CFG rc=$?"
    python - <<PY
import json
d=json.loads(open("gpurun_out/${TAG}_${name}_${CFG}.json").read().strip().splitlines()[-1])
print("$name-$CFG value", round(d["value"]), {k: round(v["ms_per_step"], 4) for k, v in d.get("kernels", {}).items()})
PY
  done
}
run new VS_NOP=1
run noskip VS_ST_SKIP=0
[ -f _variants/libvs_base.so ] && run base VS_TRACKER_LIB=$PWD/_variants/libvs_base.so
CMD="python bench.py --steps 4 --warmup 3 --device-only --repeats 1"
ncu --set full --clock-control none --import-source on -k "regex:^shi_tomasi_kernel" -s 4 -c 1 -o gpurun_out/${TAG}_st_new -f $CMD > gpurun_out/${TAG}_ncu_new.log 2>&1; echo "ncu rc=$?"
