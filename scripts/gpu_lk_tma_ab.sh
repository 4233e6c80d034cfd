# usage: bash scripts/gpu_lk_tma_ab.sh <tag>  — LK tile staging A/B: cp.async (shipped) vs cp.async.bulk.tensor (VS_LK_TMA=1):
# the LK / sequence / golden GPU tests with the TMA variant, then C2 / C1 device-leg bench lines of both
TAG=$1
mkdir -p gpurun_out
VS_LK_TMA=1 timeout 600 python -m pytest tests -m gpu -x -q -k "lk or sequence or golden or ref_pin" > gpurun_out/${TAG}_pytest_tma.txt 2>&1; echo "pytest(tma) rc=$?"; tail -4 gpurun_out/${TAG}_pytest_tma.txt
run() {  # name, env...
  local name=$1; shift
  for CFG in C2 C1; do
    env "$@" timeout 300 python bench.py --config $CFG --no-cpu --no-c1 --latency-frames 100 > gpurun_out/${TAG}_${name}_${CFG}.json 2> gpurun_out/${TAG}_${name}_${CFG}.err; echo "$name $CFG rc=$?"
    python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/${TAG}_${name}_${CFG}.json").read().strip().splitlines()[-1])
    print("$name-$CFG value", round(d["value"]), "ms/step", round(d["ms_per_step"], 4), "lk", round(d["kernels"]["lk"]["ms_per_step"], 4), "lat p50", round(d["latency_us"]["p50"], 1))
except Exception as e:
    print("$name-$CFG failed", e)
PY
  done
}
run cpasync VS_LK_TMA=0
run tma VS_LK_TMA=1
