# usage: ab.sh spp lib1 lib2 ...  (runs on the GPU box from repo root; PT_SW_UNFUSED etc. pass through)
spp=${1:-64}; shift
for rep in 1 2; do
for lib in "$@"; do
  echo "== $lib $(PT_LIB=$PWD/pathtrace_b200/lib/$lib timeout 100 python tools/quick_bench.py 1920 1080 $spp | tail -1 | cut -c60-130)"
done
done
