set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out/r3n
O=gpurun_out/r3n/cs.txt
for rep in 1 2; do
for v in product cs1 cs2; do
  if [ $v = product ]; then unset MVOLPS_LIB; else export MVOLPS_LIB=build/variants/$v.so; fi
  echo "== $v batch 592" >> $O; timeout 120 python profiles/prof_batch.py 592 256 512 >> $O 2>&1
done
done
unset MVOLPS_LIB
echo "== product cfg2" >> $O; timeout 120 python profiles/prof_solve.py --iters 100000 --reps 3 2>&1 | tail -1 >> $O
grep "==\|LPs\|pivots" $O | cut -c1-150
