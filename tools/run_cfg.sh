# usage: bash tools/run_cfg.sh <tag> — one GPU: gpu tests, configs 1-4 through the host engine, per-pass times + launch list + full capture of config 3
TAG=${1:-c}
mkdir -p gpurun_out
if [ "$2" != "nopytest" ]; then
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_$TAG.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_$TAG.log
fi
timeout 600 python tools/bench_configs.py --frames 300 > gpurun_out/configs_$TAG.json 2> gpurun_out/configs_$TAG.err; echo "configs rc=$?"; cat gpurun_out/configs_$TAG.json | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print(' config',d['config'],'fps',round(d['frames_per_s'],1),'tris',d['submitted_triangles'],'objs',d['visible_objects'])"
timeout 300 python tools/profile_config.py 3 4 > gpurun_out/passes_c3_$TAG.txt 2>&1; tail -4 gpurun_out/passes_c3_$TAG.txt
CMD="python tools/profile_config.py 3 2"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches_c3_$TAG.csv $CMD > /dev/null 2>&1; echo "list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"geometry_cutout_kernel|large_raster_kernel|shade_kernel|geometry_kernel" -s 5 -c 7 -f -o gpurun_out/prof_c3_$TAG $CMD > gpurun_out/ncu_full_c3_$TAG.log 2>&1; echo "full rc=$?"
python - <<PY
import csv
rows=[r for r in csv.reader(open('gpurun_out/launches_c3_$TAG.csv')) if len(r)>5]
H=rows[0]; ki=H.index('Kernel Name'); vi=H.index('Metric Value')
for r in rows[-12:]:
    try: print(f"{float(r[vi].replace(',',''))/1000:10.1f} us  {r[ki][:70]}")
    except: pass
PY
