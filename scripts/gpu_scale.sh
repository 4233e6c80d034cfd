# usage (inside gpurun --gpus N): bash scripts/gpu_scale.sh <N> <tag>
# C5 as BASELINE.json writes it (64 C1 streams in total, strong scaling) and the default C2 workload (512 streams per GPU,
# weak scaling) on N GPUs of one box, one rank per GPU, plus the box's PCIe / NUMA topology for the e2e (host-frame) numbers.
N=$1; TAG=$2
mkdir -p gpurun_out
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541"
[ "$N" = 1 ] && RUN="python"
$RUN bench.py --gpus $N --config C5 --no-cpu > gpurun_out/${TAG}_bench_c5_n$N.json 2> gpurun_out/${TAG}_bench_c5_n$N.err; echo "c5 n=$N rc=$?"
$RUN bench.py --gpus $N --no-cpu --no-c1 > gpurun_out/${TAG}_bench_c2_n$N.json 2> gpurun_out/${TAG}_bench_c2_n$N.err; echo "c2 n=$N rc=$?"
python scripts/h2d_linkmap.py $N > gpurun_out/${TAG}_linkmap_n$N.json 2> gpurun_out/${TAG}_linkmap_n$N.err; echo "linkmap rc=$?"; cat gpurun_out/${TAG}_linkmap_n$N.json
{ nvidia-smi topo -m; echo; lscpu | grep -E "^CPU\(s\)|Model name|Socket|NUMA"; echo; lspci -tv 2>/dev/null | head -80; echo; free -g | head -2; } > gpurun_out/${TAG}_topo_n$N.txt 2>&1
python - <<PY
import json
for c in ("c5","c2"):
    try:
        d=json.loads(open("gpurun_out/${TAG}_bench_%s_n$N.json"%c).read().strip().splitlines()[-1])
        e=d["e2e"]
        print(c,"n=$N value",round(d["value"]),"ms/step",round(d["ms_per_step"],4),"e2e",round(e["value"]),"link all",round(e["h2d_link_gbs_all_ranks_sum"],1),"solo",[round(x,1) for x in e["h2d_link_gbs_solo_per_gpu"]],"frac",round(e["frac_of_link_bound"],3),"gather_ms",e["gather_ms"],"frame_frac",round(d["frame_hbm_frac"],4))
    except Exception as ex:
        print(c,"failed",ex)
PY
