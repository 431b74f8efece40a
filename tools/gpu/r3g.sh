set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out/r3g
(time timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q) > gpurun_out/r3g/noalloc_alone.log 2>&1; echo "alone rc=$?"; tail -n 3 gpurun_out/r3g/noalloc_alone.log
(time timeout 1500 python -m pytest tests -m gpu -x -q) > gpurun_out/r3g/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 4 gpurun_out/r3g/pytest_gpu.log
(time timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')") > gpurun_out/r3g/smoke.log 2>&1; echo "smoke rc=$?"; tail -n 2 gpurun_out/r3g/smoke.log
(time timeout 600 python bench.py --impl reference) > gpurun_out/r3g/bench_ref.log 2>&1; echo "ref rc=$?"; tail -c 300 gpurun_out/r3g/bench_ref.log
(time timeout 600 python bench.py) > gpurun_out/r3g/bench_n1.log 2>&1; echo "bench rc=$?"; tail -c 300 gpurun_out/r3g/bench_n1.log
