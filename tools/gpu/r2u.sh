set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out/r2u
O=gpurun_out/r2u/qp.txt
for rep in 1 2; do
for v in head2 pf4serial product; do
  if [ $v = product ]; then unset MVOLPS_LIB; else export MVOLPS_LIB=build/variants/$v.so; fi
  echo "== $v batch 592" >> $O; timeout 120 python profiles/prof_batch.py 592 256 512 >> $O 2>&1
  echo "== $v cfg2 (13403 pivots)" >> $O; timeout 120 python profiles/prof_solve.py --iters 100000 --reps 3 2>&1 | tail -1 >> $O
done
done
unset MVOLPS_LIB
echo "== product batch 4096" >> $O; timeout 200 python profiles/prof_batch.py 4096 256 512 >> $O 2>&1
grep "==\|LPs\|pivots" $O | cut -c1-150
(time timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_res.py -m gpu -x -q) > gpurun_out/r2u/pytest.log 2>&1; tail -n 4 gpurun_out/r2u/pytest.log
