set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out/r2d
nvidia-smi -L
(time python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tests/dist_parity_nccl.py) > gpurun_out/r2d/parity.log 2>&1; echo "parity rc=$?"; tail -c 1500 gpurun_out/r2d/parity.log
# the C++ launcher: one process per GPU forked by the CLI itself, NCCL id through pipes
./mvolps_b200/bin/mvolps_b200 -f examples/mdkp_12x4_s1.lp > gpurun_out/r2d/cli_serial.txt 2>&1
./mvolps_b200/bin/mvolps_b200 -f examples/mdkp_12x4_s1.lp -np 2 -dm 1 > gpurun_out/r2d/cli_np2_sync.txt 2>&1; echo "cli sync rc=$?"
./mvolps_b200/bin/mvolps_b200 -f examples/mdkp_12x4_s1.lp -np 2 -dm 0 > gpurun_out/r2d/cli_np2_steal.txt 2>&1; echo "cli steal rc=$?"
tail -3 gpurun_out/r2d/cli_serial.txt gpurun_out/r2d/cli_np2_sync.txt gpurun_out/r2d/cli_np2_steal.txt
diff <(grep -v "^\[" gpurun_out/r2d/cli_serial.txt) <(grep -v "^\[" gpurun_out/r2d/cli_np2_sync.txt) > gpurun_out/r2d/cli_diff.txt; echo "tree diff lines: $(wc -l < gpurun_out/r2d/cli_diff.txt)"
(time python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 5 --warmup 3) > gpurun_out/r2d/bench_n2.log 2>&1; echo "bench rc=$?"; tail -c 2500 gpurun_out/r2d/bench_n2.log
(time python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 2 --workload bnb --dist-mode 1 --steps 2 --warmup 1) > gpurun_out/r2d/bnb_sync_n2.log 2>&1; tail -c 1200 gpurun_out/r2d/bnb_sync_n2.log
