set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out/r2c
(time python -m pytest tests -m gpu -x -q) > gpurun_out/r2c/pytest.log 2>&1; tail -8 gpurun_out/r2c/pytest.log
MVOLPS_BNB_TIMING=1 python bench.py --workload bnb --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/r2c/bnb.log 2>&1; grep "bnb timing" gpurun_out/r2c/bnb.log | tail -2; tail -c 400 gpurun_out/r2c/bnb.log
MVOLPS_BNB_TIMING=1 python bench.py --workload bnb --steps 1 --warmup 1 --no-cpu-baseline --cut-mode 2 --nodes 20000 > gpurun_out/r2c/bnb_cm2.log 2>&1; grep "bnb timing" gpurun_out/r2c/bnb_cm2.log | tail -1; tail -c 300 gpurun_out/r2c/bnb_cm2.log
python profiles/prof_batch.py 2048 50 200 > gpurun_out/r2c/prof_small.log 2>&1; cat gpurun_out/r2c/prof_small.log
ncu --set full --clock-control none --import-source on -k regex:simplex_res_kernel -c 1 -o gpurun_out/r2c/res_small python profiles/prof_batch.py 2048 50 200 > gpurun_out/r2c/ncu_res.log 2>&1; tail -3 gpurun_out/r2c/ncu_res.log
ls -la gpurun_out/r2c
