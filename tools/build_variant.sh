# usage: bash tools/build_variant.sh <name> [git-rev] [extra nvcc flags...]
# Builds exp/libvkr_<name>.so for tools/run_ab.sh from the working tree, or from the sources of a git revision
# (e.g. `bash tools/build_variant.sh morton 5312747` = the library before the k-d cluster partition), optionally with
# extra -D flags (VKR_GEOM_MIN_CTAS, VKR_SHADE_MIN_CTAS, VKR_GEOM_WARPS, VKR_OCC_MAX_PIXELS).
set -e
NAME=$1; REV=${2:-}; shift; [ -n "$REV" ] && shift
ROOT=$(cd "$(dirname "$0")/.." && pwd)
mkdir -p "$ROOT/exp"
SRC="$ROOT"
if [ -n "$REV" ] && [ "$REV" != "-" ]; then
  SRC=$(mktemp -d)
  git -C "$ROOT" archive "$REV" vk_renderer_b200/csrc include | tar -x -C "$SRC"      # (leaves index and working tree alone)
fi
nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo "$@" -Xcompiler -fPIC,-ffp-contract=off -shared \
  -o "$ROOT/exp/libvkr_$NAME.so" "$SRC/vk_renderer_b200/csrc/vkr_api.cu" -lnccl
echo "built exp/libvkr_$NAME.so from ${REV:-the working tree} $*"
