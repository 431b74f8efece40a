set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out/r3m
(time timeout 1500 python -m pytest tests -m gpu -x -q) > gpurun_out/r3m/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 4 gpurun_out/r3m/pytest_gpu.log
(MVOLPS_BNB_TIMING=1 timeout 300 python bench.py --workload bnb --cut-mode 2 --nodes 20000 --steps 3 --warmup 1 --no-cpu-baseline) > gpurun_out/r3m/bnb_cuts.log 2>&1; grep "bnb timing" gpurun_out/r3m/bnb_cuts.log | tail -1 | cut -c1-320; grep -o '"value": [0-9.]*' gpurun_out/r3m/bnb_cuts.log | head -1
