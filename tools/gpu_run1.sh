set -x
cd $GRAFT_REPO_ROOT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv
timeout 1500 python -m pytest tests -m gpu -q -s -rA -p no:cacheprovider > gpurun_out/r2_gpu_tests_1.txt 2>&1; echo "pytest rc=$?"
tail -40 gpurun_out/r2_gpu_tests_1.txt
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/r2_bench_1.json 2> gpurun_out/r2_bench_1.err; echo "bench rc=$?"
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_smoke_1.txt 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/r2_smoke_1.txt
timeout 120 tools/red_bench > gpurun_out/r2_red_bench.txt 2>&1
timeout 120 tools/tma_gather_bench 1 > gpurun_out/r2_tma_gather_box1.txt 2>&1; echo "tma1 rc=$?"
timeout 120 tools/tma_gather_bench 4 > gpurun_out/r2_tma_gather_box4.txt 2>&1; echo "tma4 rc=$?"
cat gpurun_out/r2_tma_gather_box1.txt gpurun_out/r2_tma_gather_box4.txt
