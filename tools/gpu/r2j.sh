set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out/r2j
(time timeout 1500 python -m pytest tests -m gpu -x -q) > gpurun_out/r2j/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 5 gpurun_out/r2j/pytest_gpu.log
