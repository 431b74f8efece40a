set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out/r2s
O=gpurun_out/r2s/pf.txt
for rep in 1 2; do
for v in product pf2 pf4 pf8; do
  if [ $v = product ]; then unset MVOLPS_LIB; else export MVOLPS_LIB=build/variants/$v.so; fi
  echo "== $v batch 592" >> $O; timeout 120 python profiles/prof_batch.py 592 256 512 >> $O 2>&1
  echo "== $v cfg2 (13403 pivots)" >> $O; timeout 120 python profiles/prof_solve.py --iters 100000 --reps 3 2>&1 | tail -1 >> $O
done
done
export MVOLPS_LIB=build/variants/pf4.so
echo "== pf4 batch 4096" >> $O; timeout 200 python profiles/prof_batch.py 4096 256 512 >> $O 2>&1
echo "== pf4 batch 592 K=14" >> $O; MVOLPS_LZ_K=14 timeout 200 python profiles/prof_batch.py 592 256 512 >> $O 2>&1
echo "== pf4 batch 592 K=10 (ring)" >> $O; MVOLPS_LZ_K=10 timeout 200 python profiles/prof_batch.py 592 256 512 >> $O 2>&1
echo "== pf4 batch 592 K=8 (ring)" >> $O; MVOLPS_LZ_K=8 timeout 200 python profiles/prof_batch.py 592 256 512 >> $O 2>&1
grep "==\|LPs\|pivots" $O | cut -c1-150
