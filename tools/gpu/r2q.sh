set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out/r2q
O=gpurun_out/r2q/ab2.txt
for rep in 1 2; do
for v in head product qserial k12; do
  if [ $v = product ]; then unset MVOLPS_LIB; else export MVOLPS_LIB=build/variants/$v.so; fi
  echo "== $v batch 592" >> $O; timeout 120 python profiles/prof_batch.py 592 256 512 >> $O 2>&1
  echo "== $v cfg2 (13403 pivots)" >> $O; timeout 120 python profiles/prof_solve.py --iters 100000 --reps 3 2>&1 | tail -1 >> $O
done
done
export MVOLPS_LIB=build/variants/k12.so
echo "== k12 K=10 batch 592" >> $O; MVOLPS_LZ_K=10 timeout 120 python profiles/prof_batch.py 592 256 512 >> $O 2>&1
echo "== k12 K=8 batch 592 (ring)" >> $O; MVOLPS_LZ_K=8 timeout 120 python profiles/prof_batch.py 592 256 512 >> $O 2>&1
echo "== k12 K=12 batch 4096" >> $O; timeout 200 python profiles/prof_batch.py 4096 256 512 >> $O 2>&1
unset MVOLPS_LIB
echo "== product no ring (peeled register path) batch 592" >> $O; MVOLPS_LZ_RING=0 timeout 120 python profiles/prof_batch.py 592 256 512 >> $O 2>&1
echo "== product batch 4096" >> $O; timeout 200 python profiles/prof_batch.py 4096 256 512 >> $O 2>&1
cat $O
