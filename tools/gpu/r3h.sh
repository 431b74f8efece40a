set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out/r3h
(time timeout 300 python bench.py --workload batch --lps 4096 --steps 3 --warmup 3) > gpurun_out/r3h/batch_4096.log 2>&1; tail -c 900 gpurun_out/r3h/batch_4096.log
(time timeout 300 python bench.py --workload batch --lps 8192 --rows 50 --cols 200 --steps 3 --warmup 3) > gpurun_out/r3h/batch_small.log 2>&1; tail -c 600 gpurun_out/r3h/batch_small.log
(time timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/r3h/launches_c3.csv python bench.py --workload bnb --cols 512 --rows 256 --batch-width 4096 --nodes 8192 --steps 1 --warmup 1 --no-cpu-baseline) > gpurun_out/r3h/ncu_c3.log 2>&1; echo "ncu rc=$?"
python profiles/summarize_launches.py gpurun_out/r3h/launches_c3.csv > gpurun_out/r3h/launches_c3_summary.txt; cat gpurun_out/r3h/launches_c3_summary.txt
