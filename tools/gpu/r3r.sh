set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out/r3r
(time timeout 1500 python -m pytest tests -m gpu -x -q) > gpurun_out/r3r/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 4 gpurun_out/r3r/pytest_gpu.log
(time timeout 600 python bench.py --steps 5 --warmup 3) > gpurun_out/r3r/bench_n1.log 2>&1; echo "bench rc=$?"; tail -c 300 gpurun_out/r3r/bench_n1.log
