set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out/r3a
(time timeout 1500 python -m pytest tests -m gpu -x -q) > gpurun_out/r3a/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 8 gpurun_out/r3a/pytest_gpu.log
(MVOLPS_BNB_TIMING=1 timeout 300 python bench.py --workload bnb --steps 3 --warmup 1 --no-cpu-baseline) > gpurun_out/r3a/bnb_pipe.log 2>&1; grep "bnb timing" gpurun_out/r3a/bnb_pipe.log | tail -2; tail -c 400 gpurun_out/r3a/bnb_pipe.log
(MVOLPS_BNB_NO_PIPELINE=1 MVOLPS_BNB_TIMING=1 timeout 300 python bench.py --workload bnb --steps 3 --warmup 1 --no-cpu-baseline) > gpurun_out/r3a/bnb_nopipe.log 2>&1; grep "bnb timing" gpurun_out/r3a/bnb_nopipe.log | tail -2; tail -c 400 gpurun_out/r3a/bnb_nopipe.log
(MVOLPS_BNB_TIMING=1 timeout 300 python bench.py --workload bnb --cut-mode 2 --nodes 20000 --steps 3 --warmup 1 --no-cpu-baseline) > gpurun_out/r3a/bnb_cuts.log 2>&1; grep "bnb timing" gpurun_out/r3a/bnb_cuts.log | tail -1; tail -c 300 gpurun_out/r3a/bnb_cuts.log
