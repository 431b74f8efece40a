set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out/r2o
timeout 300 ./build/micro/micro_streams > gpurun_out/r2o/micro_streams.txt 2>&1
cat gpurun_out/r2o/micro_streams.txt
