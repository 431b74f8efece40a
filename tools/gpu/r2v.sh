set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out/r2v
MVOLPS_LIB=build/variants/prof.so timeout 120 python profiles/prof_phases.py 1024 2048 > gpurun_out/r2v/phases_cfg2.txt 2>&1
cat gpurun_out/r2v/phases_cfg2.txt
(MVOLPS_BNB_TIMING=1 timeout 300 python bench.py --workload bnb --steps 3 --warmup 1) > gpurun_out/r2v/bnb.log 2>&1; grep "bnb timing" gpurun_out/r2v/bnb.log | tail -3; tail -c 900 gpurun_out/r2v/bnb.log
