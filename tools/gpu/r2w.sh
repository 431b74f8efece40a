set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out/r2w
(time timeout 900 python -m pytest tests/test_gpu_bnb.py -m gpu -x -q) > gpurun_out/r2w/pytest_bnb.log 2>&1; tail -n 4 gpurun_out/r2w/pytest_bnb.log
(MVOLPS_BNB_TIMING=1 timeout 300 python bench.py --workload bnb --cut-mode 2 --nodes 20000 --steps 3 --warmup 1 --no-cpu-baseline) > gpurun_out/r2w/bnb_cuts.log 2>&1; grep "bnb timing" gpurun_out/r2w/bnb_cuts.log | tail -3; tail -c 600 gpurun_out/r2w/bnb_cuts.log
