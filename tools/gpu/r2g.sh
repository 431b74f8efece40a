set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out/r2g
(time timeout 400 python -m pytest tests -m gpu -x -q) > gpurun_out/r2g/pytest.log 2>&1; tail -6 gpurun_out/r2g/pytest.log
python profiles/prof_batch.py 2048 50 200 > gpurun_out/r2g/prof_small.log 2>&1; cat gpurun_out/r2g/prof_small.log
MVOLPS_CRES=1 timeout 120 python profiles/prof_batch.py 30 256 512 > gpurun_out/r2g/prof_cres30.log 2>&1; cat gpurun_out/r2g/prof_cres30.log
MVOLPS_BNB_TIMING=1 python bench.py --workload bnb --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/r2g/bnb.log 2>&1; grep "bnb timing" gpurun_out/r2g/bnb.log | cut -c1-300; tail -c 300 gpurun_out/r2g/bnb.log
(time python bench.py --steps 5 --warmup 3) > gpurun_out/r2g/bench.log 2>&1; echo "bench rc=$?"; tail -c 1500 gpurun_out/r2g/bench.log
timeout 300 ncu --set full --clock-control none --import-source on -k regex:simplex_res_kernel -c 1 -o gpurun_out/r2g/res_small python profiles/prof_batch.py 2048 50 200 > gpurun_out/r2g/ncu_res.log 2>&1
timeout 300 ncu --set full --clock-control none -k regex:simplex_kernel -c 1 -o gpurun_out/r2g/gen_phase1 python profiles/prof_phase1.py 256 512 > gpurun_out/r2g/ncu_gen.log 2>&1
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2g/launches_bench.csv python bench.py --headline-only --no-cpu-baseline --steps 2 --warmup 3 > gpurun_out/r2g/ncu_launches.log 2>&1
