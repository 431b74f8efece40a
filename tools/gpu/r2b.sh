set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out/r2b
(time python -m pytest tests/test_gpu_res.py -x -q) > gpurun_out/r2b/pytest_res.log 2>&1; tail -25 gpurun_out/r2b/pytest_res.log
(time python -m pytest tests -m gpu -x -q) > gpurun_out/r2b/pytest.log 2>&1; tail -8 gpurun_out/r2b/pytest.log
python bench.py --workload dense --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r2b/dense.log 2>&1; tail -c 1200 gpurun_out/r2b/dense.log
MVOLPS_BNB_TIMING=1 python bench.py --workload bnb --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/r2b/bnb.log 2>&1; grep "bnb timing" gpurun_out/r2b/bnb.log | tail -2; tail -c 600 gpurun_out/r2b/bnb.log
MVOLPS_NO_RES=1 MVOLPS_BNB_TIMING=1 python bench.py --workload bnb --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/r2b/bnb_nores.log 2>&1; grep "bnb timing" gpurun_out/r2b/bnb_nores.log | tail -2
python bench.py --workload batch --rows 50 --cols 200 --lps 8192 --steps 3 --warmup 3 > gpurun_out/r2b/batch_small_res.log 2>&1; tail -c 900 gpurun_out/r2b/batch_small_res.log
MVOLPS_NO_RES=1 python bench.py --workload batch --rows 50 --cols 200 --lps 8192 --steps 3 --warmup 3 > gpurun_out/r2b/batch_small_nores.log 2>&1; tail -c 900 gpurun_out/r2b/batch_small_nores.log
