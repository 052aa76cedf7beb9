# usage: ab_trav.sh spp trav1 trav2 ...   (development helper: same-box A/B of PT_TRAVERSAL_* values on C2's frame)
spp=${1:-64}; shift
for rep in 1 2; do
for t in "$@"; do
  echo "== $t $(timeout 100 python tools/quick_bench.py 1920 1080 $spp $t | tail -1 | cut -c40-150)"
done
done
