# usage: bash tools/run_ncu_cfg.sh <tag> <config> [kernel regex]
TAG=${1:-c3}; CFG=${2:-3}; KRE=${3:-"large_raster_kernel|geometry_cutout_kernel"}
mkdir -p gpurun_out
CMD="python tools/profile_config.py $CFG 2"
$CMD > gpurun_out/plain_$TAG.log 2>&1 || exit 1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches_$TAG.csv $CMD > /dev/null 2>&1; echo "list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"$KRE" -s 2 -c 3 -f -o gpurun_out/prof_$TAG $CMD > gpurun_out/ncu_full_$TAG.log 2>&1; echo "full rc=$?"
python - <<PY
import csv
rows=[r for r in csv.reader(open('gpurun_out/launches_$TAG.csv')) if len(r)>5]
H=rows[0]; ki=H.index('Kernel Name'); vi=H.index('Metric Value')
for r in rows[-14:]:
    try: print(f"{float(r[vi].replace(',',''))/1000:10.1f} us  {r[ki][:70]}")
    except: pass
PY
