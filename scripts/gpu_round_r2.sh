# usage: bash scripts/gpu_round_r2.sh <tag>   — one round state's evidence: GPU test log, bench lines (C2 default with the CPU
# baseline and the C1 sub-record, C1, C3, C4, reference arm), then scripts/gpu_profile_r2.sh (launch list of the device pass
# alone + ncu --set full of every hot kernel), then the single-stream latency probe.  Everything lands in gpurun_out/<tag>_*.
TAG=$1
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/${TAG}_pytest_gpu.txt 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/${TAG}_pytest_gpu.txt
python bench.py > gpurun_out/${TAG}_bench_c2.json 2> gpurun_out/${TAG}_bench_c2.err; echo "bench c2 rc=$?"
python bench.py --config C1 --no-cpu > gpurun_out/${TAG}_bench_c1.json 2> gpurun_out/${TAG}_bench_c1.err; echo "bench c1 rc=$?"
python bench.py --config C3 --streams 256 --no-cpu > gpurun_out/${TAG}_bench_c3.json 2> gpurun_out/${TAG}_bench_c3.err; echo "bench c3 rc=$?"
python bench.py --config C4 --streams 128 --no-cpu > gpurun_out/${TAG}_bench_c4.json 2> gpurun_out/${TAG}_bench_c4.err; echo "bench c4 rc=$?"
python bench.py --impl reference --steps 5 --warmup 2 > gpurun_out/${TAG}_bench_ref.json 2> gpurun_out/${TAG}_bench_ref.err; echo "ref rc=$?"
bash scripts/gpu_profile_r2.sh ${TAG}
python scripts/latency_probe.py C2 300 > gpurun_out/${TAG}_latency.txt 2>&1; python scripts/latency_probe.py C1 300 >> gpurun_out/${TAG}_latency.txt 2>&1
python scripts/latency_probe.py C2 300 off >> gpurun_out/${TAG}_latency.txt 2>&1; tail -6 gpurun_out/${TAG}_latency.txt
