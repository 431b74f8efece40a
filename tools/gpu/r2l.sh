set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out/r2l
(time timeout 900 python -m pytest tests/test_gpu_shard.py -m gpu -x -q) > gpurun_out/r2l/pytest_shard.log 2>&1; echo "shard rc=$?"; tail -n 15 gpurun_out/r2l/pytest_shard.log
bash tools/gpu/r2k.sh
