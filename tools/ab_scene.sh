# usage: ab_scene.sh scene W H spp lib1 lib2 ...   (development helper: same-box A/B of library builds on one scene)
scene=$1; W=$2; H=$3; spp=$4; shift 4
for lib in "$@"; do
  echo "== $lib $(PT_LIB=$PWD/pathtrace_b200/lib/$lib timeout 300 python tools/quick_bench.py $W $H $spp auto 0 $scene | tail -1 | cut -c40-200)"
done
