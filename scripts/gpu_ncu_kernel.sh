# usage: bash scripts/gpu_ncu_kernel.sh <kernel-regex> <out-name> [skip] [count] [extra bench args]
mkdir -p gpurun_out
K=$1; OUT=$2; SKIP=${3:-4}; CNT=${4:-2}; shift 4
CMD="python bench.py --steps 3 --warmup 3 --no-cpu --latency-frames 1 $@"
$CMD > gpurun_out/plain_$OUT.log 2>&1 && ncu --set full --clock-control none --import-source on -k "regex:$K" -s $SKIP -c $CNT -o gpurun_out/$OUT -f $CMD > gpurun_out/ncu_$OUT.log 2>&1
echo "rc=$?"; tail -3 gpurun_out/ncu_$OUT.log
