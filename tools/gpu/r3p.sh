set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out/r3p
(time timeout 900 python -m pytest tests/test_gpu_res.py tests/test_gpu_bnb.py -m gpu -x -q) > gpurun_out/r3p/pytest.log 2>&1; echo "rc=$?"; tail -n 4 gpurun_out/r3p/pytest.log
