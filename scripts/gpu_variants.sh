# usage: bash scripts/gpu_variants.sh <tag> <pytest -k expr> <variant libs...>  — GPU tests (-k expr) on the tree's library, then C2 / C1
# device-leg bench lines for the tree's library and every _variants/libvs_<name>.so named
TAG=$1; KEXPR=$2; shift 2
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "$KEXPR" > gpurun_out/${TAG}_pytest_gpu.txt 2>&1; echo "pytest rc=$?"; tail -25 gpurun_out/${TAG}_pytest_gpu.txt
run() {  # name, env...
  local name=$1; shift
  for CFG in C2 C1; do
    env "$@" python bench.py --config $CFG --no-cpu --no-c1 --device-only > gpurun_out/${TAG}_${name}_${CFG}.json 2> gpurun_out/${TAG}_${name}_${CFG}.err; echo "$name $CFG rc=$?"
    python - <<PY
import json
d=json.loads(open("gpurun_out/${TAG}_${name}_${CFG}.json").read().strip().splitlines()[-1])
print("$name-$CFG value", round(d["value"]), "ms/step", round(d["ms_per_step"], 4), d.get("stages", {}).get("shi_tomasi", {}).get("ms_per_step"))
PY
  done
}
run new VS_NOP=1
for v in "$@"; do [ -f _variants/libvs_$v.so ] && run $v VS_TRACKER_LIB=$PWD/_variants/libvs_$v.so; done
