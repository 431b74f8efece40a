set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out/r3j
(time timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29561 bench.py --gpus 8 --steps 5 --warmup 3) > gpurun_out/r3j/bench_n8.log 2>&1; echo "bench8 rc=$?"; tail -c 300 gpurun_out/r3j/bench_n8.log
