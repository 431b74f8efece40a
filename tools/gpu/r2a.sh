set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out/r2a
nvidia-smi -L
(time python -m pytest tests -m gpu -x -q) > gpurun_out/r2a/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2a/pytest.log
tail -5 gpurun_out/r2a/pytest.log
(time python bench.py --steps 5 --warmup 3) > gpurun_out/r2a/bench.log 2>&1; echo "bench rc=$?" >> gpurun_out/r2a/bench.log
tail -c 3000 gpurun_out/r2a/bench.log
(time python bench.py --impl reference --steps 3 --warmup 1) > gpurun_out/r2a/bench_ref.log 2>&1
python profiles/prof_phase1.py > gpurun_out/r2a/phase1_handover.log 2>&1
MVOLPS_NO_HANDOVER=1 python profiles/prof_phase1.py > gpurun_out/r2a/phase1_nohandover.log 2>&1
cat gpurun_out/r2a/phase1_*.log
python bench.py --workload batch --steps 3 --warmup 3 > gpurun_out/r2a/batch.log 2>&1; tail -c 1500 gpurun_out/r2a/batch.log
MVOLPS_BNB_TIMING=1 python bench.py --workload bnb --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/r2a/bnb.log 2>&1; tail -c 2500 gpurun_out/r2a/bnb.log
MVOLPS_BNB_TIMING=1 python bench.py --workload bnb --steps 1 --warmup 1 --no-cpu-baseline --cut-mode 2 --nodes 20000 > gpurun_out/r2a/bnb_cm2.log 2>&1; tail -c 2500 gpurun_out/r2a/bnb_cm2.log
timeout 900 compute-sanitizer --tool memcheck python profiles/sanitize_small.py > gpurun_out/r2a/sanitize_memcheck.log 2>&1; echo "memcheck rc=$?"; tail -3 gpurun_out/r2a/sanitize_memcheck.log
timeout 1200 compute-sanitizer --tool racecheck python profiles/sanitize_small.py > gpurun_out/r2a/sanitize_racecheck.log 2>&1; echo "racecheck rc=$?"; tail -3 gpurun_out/r2a/sanitize_racecheck.log
