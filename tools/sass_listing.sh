# usage: bash tools/sass_listing.sh <tag>  — resource usage of every kernel in libvkr.so and the SASS of the hot ones, under profiles/<tag>_sass/
# (instruction text only: addresses and encodings stripped; needs no GPU)
TAG=${1:-r02}; LIB=vk_renderer_b200/libvkr.so; OUT=profiles/${TAG}_sass
mkdir -p $OUT
{ echo "# cuobjdump -res-usage $LIB ($(git rev-parse --short HEAD)$(git status --porcelain -- vk_renderer_b200/csrc | grep -q . && echo +dirty))";
  cuobjdump -res-usage $LIB 2>/dev/null | grep -A1 "Function" | grep -v "^--" | paste - - | sed 's/^ *//' | sed 's/Function \(_Z[A-Za-z0-9_]*\):/\1/' | while read f rest; do echo "$(echo $f | c++filt | cut -c1-110) | $rest"; done; } > $OUT/resources.txt
for k in _Z15geometry_kernelILb0 _Z15geometry_kernelILb1 _Z12shade_kernel _Z19large_raster_kernelILb0 _Z20expand_chunks_kernel; do
  sym=$(cuobjdump -res-usage $LIB 2>/dev/null | grep -o "${k}[A-Za-z0-9_]*" | head -1)
  [ -z "$sym" ] && continue
  name=$(echo $k | sed 's/_Z[0-9]*//; s/ILb0/_main/; s/ILb1/_shadow/')
  cuobjdump -sass -fun $sym $LIB 2>/dev/null | grep -E "^\s+/\*[0-9a-f]{4,}\*/" | sed -E 's#^\s+/\*[0-9a-f]+\*/\s+##; s#\s*/\* 0x[0-9a-f]+ \*/\s*$##; s#\s+;$# ;#' > $OUT/$name.sass
  { echo "## $name: $(wc -l < $OUT/$name.sass) instructions; mnemonic histogram (top 40)"; sed -E 's/^@!?U?P[0-9T]+ //' $OUT/$name.sass | awk '{print $1}' | sort | uniq -c | sort -rn | head -40; } >> $OUT/histogram.txt
done
ls -la $OUT
