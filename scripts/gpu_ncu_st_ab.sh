# usage: bash scripts/gpu_ncu_st_ab.sh <tag>  — ncu --set full of one steady-state shi_tomasi launch (C2, 512 streams) for the
# tree's library, the same with VS_ST_SKIP=0, and _variants/libvs_base.so
TAG=$1
mkdir -p gpurun_out
CMD="python bench.py --steps 4 --warmup 3 --device-only --repeats 1"
cap() { # name env...
  local name=$1; shift
  env "$@" $CMD > gpurun_out/${TAG}_${name}_plain.log 2>&1 || { echo "$name plain run failed"; return; }
  env "$@" ncu --set full --clock-control none --import-source on -k "regex:^shi_tomasi_kernel" -s 4 -c 1 -o gpurun_out/${TAG}_st_${name} -f $CMD > gpurun_out/${TAG}_ncu_${name}.log 2>&1; echo "$name rc=$?"
}
cap new VS_NOP=1
cap noskip VS_ST_SKIP=0
[ -f _variants/libvs_base.so ] && cap base VS_TRACKER_LIB=$PWD/_variants/libvs_base.so
