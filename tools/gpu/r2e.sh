set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out/r2e
(time timeout 300 python -m pytest tests -m gpu -x -q -k "not cluster_resident") > gpurun_out/r2e/pytest.log 2>&1; tail -6 gpurun_out/r2e/pytest.log
MVOLPS_BNB_TIMING=1 python bench.py --workload bnb --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/r2e/bnb.log 2>&1; grep "bnb timing" gpurun_out/r2e/bnb.log | cut -c1-300
MVOLPS_BNB_TIMING=1 python bench.py --workload bnb --steps 1 --warmup 1 --no-cpu-baseline --cut-mode 2 --nodes 20000 > gpurun_out/r2e/bnb_cm2.log 2>&1; grep "bnb timing" gpurun_out/r2e/bnb_cm2.log | tail -1 | cut -c1-300
python profiles/prof_batch.py 2048 50 200 > gpurun_out/r2e/prof_small.log 2>&1; cat gpurun_out/r2e/prof_small.log
# streaming batch: occupancy / depth sweep of the lazy loop (cluster-resident off)
for cfg in "0 0" "128 8" "128 6" "256 6" "128 5"; do set -- $cfg; MVOLPS_CRES=0 MVOLPS_PS_NT=$1 MVOLPS_LZ_K=$2 python profiles/prof_batch.py 592 256 512 2>&1 | tail -1 | sed "s/^/nt=$1 K=$2: /"; done > gpurun_out/r2e/stream_sweep.log 2>&1; cat gpurun_out/r2e/stream_sweep.log
# cluster-resident kernel, under a timeout of its own
(time timeout 240 python -m pytest tests/test_gpu_res.py -x -q -k "cluster_resident") > gpurun_out/r2e/pytest_cres.log 2>&1; echo "cres rc=$?"; tail -30 gpurun_out/r2e/pytest_cres.log | cut -c1-300
timeout 120 python profiles/prof_batch.py 592 256 512 > gpurun_out/r2e/prof_cres.log 2>&1; cat gpurun_out/r2e/prof_cres.log
