# usage: bash tools/run_ab.sh <tag> "<variant list>"  — A/B of prebuilt library variants exp/libvkr_<v>.so on the bench workload
TAG=${1:-ab}; VS=${2:-"c4 c5"}
mkdir -p gpurun_out
cp vk_renderer_b200/libvkr.so /tmp/libvkr_orig.so
for v in $VS; do
  cp exp/libvkr_$v.so vk_renderer_b200/libvkr.so
  for rep in 1 2; do
    timeout 300 python bench.py --steps 100 --warmup 5 --no-cpu-baseline --no-views --no-scene-update > gpurun_out/ab_${TAG}_$v.json 2> gpurun_out/ab_${TAG}_$v.err; rc=$?
    python -c "
import json;d=json.load(open('gpurun_out/ab_${TAG}_$v.json'));p=d['pass_ms'];print('$v rc=$rc fps',round(d['value'],1),'shadow',round(p['shadow'],3),'geom_k',round(p['geom_kernel'],3),'shade',round(p['shade'],3),'clear',round(p['clear'],3))"
  done
done
cp /tmp/libvkr_orig.so vk_renderer_b200/libvkr.so
