set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out/r2r
(time timeout 1500 python -m pytest tests -m gpu -x -q) > gpurun_out/r2r/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 6 gpurun_out/r2r/pytest_gpu.log
O=gpurun_out/r2r/ab3.txt
for rep in 1 2; do
for v in head product; do
  if [ $v = product ]; then unset MVOLPS_LIB; else export MVOLPS_LIB=build/variants/$v.so; fi
  echo "== $v batch 592" >> $O; timeout 120 python profiles/prof_batch.py 592 256 512 >> $O 2>&1
  echo "== $v cfg2 (13403 pivots)" >> $O; timeout 120 python profiles/prof_solve.py --iters 100000 --reps 3 2>&1 | tail -1 >> $O
done
done
unset MVOLPS_LIB
echo "== product batch 4096" >> $O; timeout 200 python profiles/prof_batch.py 4096 256 512 >> $O 2>&1
echo "== product batch 592 K=14" >> $O; MVOLPS_LZ_K=14 timeout 200 python profiles/prof_batch.py 592 256 512 >> $O 2>&1
echo "== product batch 592 K=10" >> $O; MVOLPS_LZ_K=10 timeout 200 python profiles/prof_batch.py 592 256 512 >> $O 2>&1
echo "== product batch 8192 x 50x200 (resident)" >> $O; timeout 200 python profiles/prof_batch.py 8192 50 200 >> $O 2>&1
echo "== prof batch 592" >> $O; MVOLPS_LIB=build/variants/prof.so timeout 120 python profiles/prof_batch.py 592 256 512 >> $O 2>&1
grep "==\|LPs\|pivots\|us)" $O | cut -c1-150
(time timeout 600 python bench.py --steps 5 --warmup 3) > gpurun_out/r2r/bench_n1.log 2>&1; echo "bench rc=$?"; tail -c 300 gpurun_out/r2r/bench_n1.log
