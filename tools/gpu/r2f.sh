set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out/r2f
(time timeout 400 python -m pytest tests -m gpu -x -q) > gpurun_out/r2f/pytest.log 2>&1; tail -6 gpurun_out/r2f/pytest.log
MVOLPS_BNB_TIMING=1 python bench.py --workload bnb --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/r2f/bnb.log 2>&1; grep "bnb timing" gpurun_out/r2f/bnb.log | cut -c1-300; tail -c 300 gpurun_out/r2f/bnb.log
MVOLPS_BNB_TIMING=1 python bench.py --workload bnb --steps 1 --warmup 1 --no-cpu-baseline --cut-mode 2 --nodes 20000 > gpurun_out/r2f/bnb_cm2.log 2>&1; grep "bnb timing" gpurun_out/r2f/bnb_cm2.log | tail -1 | cut -c1-300
python profiles/prof_batch.py 2048 50 200 > gpurun_out/r2f/prof_small.log 2>&1; cat gpurun_out/r2f/prof_small.log
MVOLPS_CRES=1 timeout 120 python profiles/prof_batch.py 30 256 512 > gpurun_out/r2f/prof_cres30.log 2>&1; cat gpurun_out/r2f/prof_cres30.log
MVOLPS_CRES=0 timeout 120 python profiles/prof_batch.py 30 256 512 > gpurun_out/r2f/prof_lazy30.log 2>&1; cat gpurun_out/r2f/prof_lazy30.log
MVOLPS_CRES=1 timeout 300 ncu --set full --clock-control none --import-source on -k regex:simplex_cres_kernel -c 1 -o gpurun_out/r2f/cres30 python profiles/prof_batch.py 30 256 512 > gpurun_out/r2f/ncu_cres.log 2>&1; tail -2 gpurun_out/r2f/ncu_cres.log
timeout 300 ncu --set full --clock-control none --import-source on -k regex:simplex_res_kernel -c 1 -o gpurun_out/r2f/res_small python profiles/prof_batch.py 2048 50 200 > gpurun_out/r2f/ncu_res.log 2>&1; tail -2 gpurun_out/r2f/ncu_res.log
