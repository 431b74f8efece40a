set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out/r2h
nvidia-smi -L | wc -l
(time timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 tests/dist_parity_nccl.py) > gpurun_out/r2h/parity_n8.log 2>&1; echo "parity rc=$?"; tail -c 400 gpurun_out/r2h/parity_n8.log
(time timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29523 bench.py --gpus 8 --steps 5 --warmup 3) > gpurun_out/r2h/bench_n8.log 2>&1; echo "bench8 rc=$?"; tail -c 400 gpurun_out/r2h/bench_n8.log
(time timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29522 bench.py --gpus 4 --steps 5 --warmup 3) > gpurun_out/r2h/bench_n4.log 2>&1; echo "bench4 rc=$?"; tail -c 400 gpurun_out/r2h/bench_n4.log
(time timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29524 bench.py --gpus 8 --workload bnb --dist-mode 1 --steps 2 --warmup 1) > gpurun_out/r2h/bnb_sync_n8.log 2>&1; tail -c 300 gpurun_out/r2h/bnb_sync_n8.log
timeout 120 ./mvolps_b200/bin/mvolps_b200 -f examples/mdkp_12x4_s1.lp -np 8 -dm 1 -s > gpurun_out/r2h/cli_np8_sync.txt 2>&1; echo "cli rc=$?"; tail -n 2 gpurun_out/r2h/cli_np8_sync.txt
