# usage (on the GPU box): bash tools/round_check.sh TAG  — the no-alloc test alone, the whole gpu suite, the bench line,
# its ncu launch list, and one ncu --set full capture of the C2 kernel (development helper; outputs under gpurun_out/)
tag=${1:-rX}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/${tag}_noalloc_alone.log 2>&1; echo "noalloc alone rc=$?"
tail -3 gpurun_out/${tag}_noalloc_alone.log
timeout 2400 python -m pytest tests -x -q -m gpu > gpurun_out/${tag}_gpu_suite.log 2>&1; echo "suite rc=$?"
tail -4 gpurun_out/${tag}_gpu_suite.log
timeout 900 python bench.py > gpurun_out/${tag}_bench_n1.json 2> gpurun_out/${tag}_bench_n1.err; echo "bench rc=$?"
cut -c1-600 gpurun_out/${tag}_bench_n1.json
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_launches.csv \
  python bench.py --steps 2 --warmup 3 --no-configs --no-cpu-baseline > gpurun_out/${tag}_ncu_launches.log 2>&1; echo "launch list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_smem_wavefront --launch-skip 1 --launch-count 1 \
  -f -o gpurun_out/${tag}_full python tools/quick_bench.py 1920 1080 1024 > gpurun_out/${tag}_ncu_full.log 2>&1; echo "ncu full rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_smem_wavefront --launch-skip 1 --launch-count 1 \
  -f -o gpurun_out/${tag}_full_64spp python tools/quick_bench.py 1920 1080 64 > gpurun_out/${tag}_ncu_full64.log 2>&1; echo "ncu full64 rc=$?"
