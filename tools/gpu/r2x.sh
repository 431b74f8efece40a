set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out/r2x
(time timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu-baseline) > gpurun_out/r2x/bench_plain.log 2>&1; echo "plain rc=$?"
(time timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 20000 --csv --log-file gpurun_out/r2x/launches_bench.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline) > gpurun_out/r2x/ncu_launches.log 2>&1; echo "ncu rc=$?"
python profiles/summarize_launches.py gpurun_out/r2x/launches_bench.csv > gpurun_out/r2x/launches_summary.txt; cat gpurun_out/r2x/launches_summary.txt
gzip -9 gpurun_out/r2x/launches_bench.csv
