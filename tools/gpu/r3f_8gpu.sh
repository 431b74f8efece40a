set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out/r3f
nproc
(time timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 8 --steps 5 --warmup 3) > gpurun_out/r3f/bench_n8.log 2>&1; echo "bench8 rc=$?"; tail -c 300 gpurun_out/r3f/bench_n8.log
(time timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29542 bench.py --gpus 8 --workload bnb --cols 512 --rows 256 --batch-width 4096 --nodes 8192 --dist-mode 1 --steps 2 --warmup 1) > gpurun_out/r3f/c3_sync_n8.log 2>&1; tail -c 300 gpurun_out/r3f/c3_sync_n8.log
(time timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29543 bench.py --gpus 4 --steps 5 --warmup 3) > gpurun_out/r3f/bench_n4.log 2>&1; echo "bench4 rc=$?"; tail -c 300 gpurun_out/r3f/bench_n4.log
