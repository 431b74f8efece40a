set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out/r3b
nproc
(time timeout 1500 python -m pytest tests -m gpu -x -q) > gpurun_out/r3b/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 4 gpurun_out/r3b/pytest_gpu.log
for T in 4 1 2 8; do
(MVOLPS_HOST_THREADS=$T MVOLPS_BNB_TIMING=1 timeout 300 python bench.py --workload bnb --steps 3 --warmup 1 --no-cpu-baseline) > gpurun_out/r3b/bnb_t$T.log 2>&1; echo "T=$T"; grep "bnb timing" gpurun_out/r3b/bnb_t$T.log | sed -n 3p | cut -c1-200; grep -o '"value": [0-9.]*' gpurun_out/r3b/bnb_t$T.log | head -1
done
(time timeout 600 python bench.py --steps 5 --warmup 3) > gpurun_out/r3b/bench_n1.log 2>&1; echo "bench rc=$?"; tail -c 300 gpurun_out/r3b/bench_n1.log
