TAG=$1
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "feature_bank or feature_manager or lk_tma" > gpurun_out/${TAG}_pytest_bank.txt 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/${TAG}_pytest_bank.txt
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
for CFG in C4:128 C2:512; do
  python bench.py --config ${CFG%%:*} --streams ${CFG##*:} --no-cpu --no-c1 --latency-frames 20 > gpurun_out/${TAG}_bank_${CFG%%:*}.json 2> gpurun_out/${TAG}_bank_${CFG%%:*}.err; echo "rc=$?"
  python - <<PY
import json
d=json.loads(open("gpurun_out/${TAG}_bank_${CFG%%:*}.json").read().strip().splitlines()[-1]); print("${CFG}", round(d["value"]), d["consumers"])
PY
done
