# usage: bash tools/run_ab_cfg.sh "<variants>" "<configs>"  — A/B of exp/libvkr_<v>.so on tools/bench_configs.py
VS=${1:-"head new"}; CF=${2:-"1,2,3,4"}
cp vk_renderer_b200/libvkr.so /tmp/libvkr_orig.so
for v in $VS; do
  cp exp/libvkr_$v.so vk_renderer_b200/libvkr.so
  echo "== $v"; python tools/bench_configs.py --frames 300 --configs $CF 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    try: d=json.loads(l); print(' config',d['config'],'fps',round(d['frames_per_s'],1))
    except Exception: print(l.strip()[:200])"
done
cp /tmp/libvkr_orig.so vk_renderer_b200/libvkr.so
