# usage: bash tools/run_step.sh <tag> "<variants for the A/B>" [nopytest]  — one box, one GPU: gpu tests on the in-tree library,
# A/B of exp/libvkr_<v>.so on the bench workload, launch list and full ncu capture of the in-tree library
TAG=${1:-x}; VS=${2:-""}
mkdir -p gpurun_out
if [ "$3" != "nopytest" ]; then
  timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_$TAG.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_$TAG.log
fi
[ -n "$VS" ] && bash tools/run_ab.sh $TAG "$VS"
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-views --no-scene-update"
timeout 600 $CMD > gpurun_out/plain_$TAG.log 2>&1 && {
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_$TAG.csv $CMD > gpurun_out/ncu_list_$TAG.log 2>&1; echo "launch list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"geometry_kernel|shade_kernel" -s 12 -c 3 -f -o gpurun_out/prof_$TAG $CMD > gpurun_out/ncu_full_$TAG.log 2>&1; echo "full capture rc=$?"
}
ls -la gpurun_out | tail -8
