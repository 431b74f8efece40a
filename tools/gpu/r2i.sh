set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out/r2i
(time timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q) > gpurun_out/r2i/noalloc_alone.log 2>&1; echo "alone rc=$?"; tail -n 3 gpurun_out/r2i/noalloc_alone.log
(time timeout 1500 python -m pytest tests -m gpu -x -q) > gpurun_out/r2i/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 5 gpurun_out/r2i/pytest_gpu.log
(time timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')") > gpurun_out/r2i/smoke.log 2>&1; echo "smoke rc=$?"; tail -n 3 gpurun_out/r2i/smoke.log
(time timeout 600 python bench.py --steps 5 --warmup 3) > gpurun_out/r2i/bench_n1.log 2>&1; echo "bench rc=$?"; tail -c 1500 gpurun_out/r2i/bench_n1.log
