# usage: bash scripts/gpu_sanitize_bank.sh <tag> — compute-sanitizer (memcheck, racecheck) over the feature bank kernels and the
# segmented Shi-Tomasi walk on small cases
TAG=$1
mkdir -p gpurun_out
which compute-sanitizer || { echo "no compute-sanitizer"; exit 0; }
timeout 900 compute-sanitizer --tool memcheck --error-exitcode 9 python -m pytest tests/test_gpu_feature_bank.py -x -q -k "errors or batch or tracker" > gpurun_out/${TAG}_memcheck_bank.txt 2>&1; echo "memcheck bank rc=$?"; grep -E "ERROR SUMMARY|passed|failed" gpurun_out/${TAG}_memcheck_bank.txt | tail -3
timeout 900 compute-sanitizer --tool racecheck --error-exitcode 9 python -m pytest tests/test_gpu_feature_bank.py -x -q -k "batch" > gpurun_out/${TAG}_racecheck_bank.txt 2>&1; echo "racecheck bank rc=$?"; grep -E "RACECHECK SUMMARY|passed|failed" gpurun_out/${TAG}_racecheck_bank.txt | tail -3
timeout 900 compute-sanitizer --tool memcheck --error-exitcode 9 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_memcheck_smoke.txt 2>&1; echo "memcheck smoke rc=$?"; grep -E "ERROR SUMMARY|smoke ok" gpurun_out/${TAG}_memcheck_smoke.txt | tail -3
