set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out/r2y
O=gpurun_out/r2y/ab4.txt
for rep in 1 2; do
for v in product peelg; do
  if [ $v = product ]; then unset MVOLPS_LIB; else export MVOLPS_LIB=build/variants/$v.so; fi
  echo "== $v cfg2 (13403 pivots)" >> $O; timeout 120 python profiles/prof_solve.py --iters 100000 --reps 3 2>&1 | tail -1 >> $O
done
done
unset MVOLPS_LIB
echo "== product batch 592" >> $O; timeout 120 python profiles/prof_batch.py 592 256 512 >> $O 2>&1
export MVOLPS_LIB=build/variants/ring6.so
for K in 11 10 12; do
echo "== ring6 K=$K batch 592" >> $O; MVOLPS_LZ_K=$K timeout 120 python profiles/prof_batch.py 592 256 512 >> $O 2>&1
done
unset MVOLPS_LIB
grep "==\|LPs\|pivots" $O | cut -c1-150
