set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out/r3i
(time timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29551 tests/dist_parity_nccl.py) > gpurun_out/r3i/parity_n2.log 2>&1; echo "parity rc=$?"; tail -c 300 gpurun_out/r3i/parity_n2.log
(time timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29552 bench.py --gpus 2 --steps 5 --warmup 3) > gpurun_out/r3i/bench_n2.log 2>&1; echo "bench2 rc=$?"; tail -c 300 gpurun_out/r3i/bench_n2.log
timeout 120 ./mvolps_b200/bin/mvolps_b200 -f examples/mdkp_12x4_s1.lp -np 2 -s > gpurun_out/r3i/cli_np2.txt 2>&1; echo "cli rc=$?"; tail -n 2 gpurun_out/r3i/cli_np2.txt
