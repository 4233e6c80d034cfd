# quick GPU check: parity tests + one bench line; usage: bash scripts/gpu_quick.sh <tag> [bench args]
mkdir -p gpurun_out
TAG=$1; shift
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15
python bench.py --no-cpu --latency-frames 50 "$@" 2>gpurun_out/bench_$TAG.err | tail -1 > gpurun_out/bench_$TAG.json
python - <<PY
import json
d=json.load(open("gpurun_out/bench_$TAG.json"))
print("frames/s", round(d["value"]), "ms/step", round(d["ms_per_step"],4), "e2e", round(d["e2e"]["value"]), "lat", d["latency_us"])
print({k:round(v["ms_per_step"],4) for k,v in d["kernels"].items()})
PY
