set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out/r2z
(time timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus 8 --steps 5 --warmup 3) > gpurun_out/r2z/bench_n8.log 2>&1; echo "bench8 rc=$?"; tail -c 300 gpurun_out/r2z/bench_n8.log
(time timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29532 bench.py --gpus 8 --workload bnb --n 512 --m 256 --batch-width 4096 --nodes 8192 --dist-mode 1 --steps 2 --warmup 1) > gpurun_out/r2z/c3_sync_n8.log 2>&1; tail -c 300 gpurun_out/r2z/c3_sync_n8.log
(time timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 5 --warmup 3) > gpurun_out/r2z/bench_n2.log 2>&1; echo "bench2 rc=$?"; tail -c 300 gpurun_out/r2z/bench_n2.log
