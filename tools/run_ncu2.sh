# usage: bash tools/run_ncu2.sh <tag> [kernel regex]  — full capture only (after a plain run)
TAG=${1:-x}
KRE=${2:-"geometry_kernel|shade_kernel"}
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/plain_$TAG.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"$KRE" -s 9 -c 3 -f -o gpurun_out/prof_$TAG $CMD > gpurun_out/ncu_full_$TAG.log 2>&1
echo "full capture rc=$?"
