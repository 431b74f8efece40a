set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out/r3q
(time timeout 1500 python -m pytest tests -m gpu -x -q) > gpurun_out/r3q/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 4 gpurun_out/r3q/pytest_gpu.log
for v in async noasync async noasync; do
  if [ $v = noasync ]; then export MVOLPS_BNB_NO_ASYNC_PULL=1; else unset MVOLPS_BNB_NO_ASYNC_PULL; fi
  (MVOLPS_BNB_TIMING=1 timeout 300 python bench.py --workload bnb --steps 3 --warmup 1 --no-cpu-baseline) > gpurun_out/r3q/bnb_$v.log 2>&1; echo "$v"; grep "bnb timing" gpurun_out/r3q/bnb_$v.log | sed -n 3p | cut -c1-200; grep -o '"value": [0-9.]*' gpurun_out/r3q/bnb_$v.log | head -1; grep -o '"ms_to_optimum": [0-9.]*' gpurun_out/r3q/bnb_$v.log
done
unset MVOLPS_BNB_NO_ASYNC_PULL
(time timeout 600 python bench.py --steps 5 --warmup 3) > gpurun_out/r3q/bench_n1.log 2>&1; echo "bench rc=$?"; tail -c 300 gpurun_out/r3q/bench_n1.log
