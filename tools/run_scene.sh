# usage: bash tools/run_scene.sh <tag> [configs]   — one box, one GPU: the device-side update_scene tests, host vs device update_scene on
# configs 6 / 4 / 1, then the whole GPU suite and the bench line
TAG=${1:-x}
CONFIGS=${2:-6,4,1}
mkdir -p gpurun_out
timeout 240 python -m pytest tests/test_scene_update.py -m gpu -x -q > gpurun_out/pytest_scene_$TAG.log 2>&1; echo "scene pytest rc=$?"; tail -15 gpurun_out/pytest_scene_$TAG.log
for mode in "" "--gpu-scene"; do
  timeout 120 python tools/bench_configs.py --configs $CONFIGS --frames 200 --loop-flags $mode >> gpurun_out/scene_configs_$TAG.json 2>> gpurun_out/scene_configs_$TAG.err; echo "configs $mode rc=$?"
done
cat gpurun_out/scene_configs_$TAG.json; tail -3 gpurun_out/scene_configs_$TAG.err
timeout 420 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_$TAG.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_$TAG.log
timeout 240 python bench.py --steps 200 --warmup 5 > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; echo "bench rc=$?"; tail -2 gpurun_out/bench_$TAG.err
python -c "
import json;d=json.load(open('gpurun_out/bench_$TAG.json'));print('fps',d['value'],'e2e',d['e2e']['value'],'roofline',d['roofline']['frac'])"
