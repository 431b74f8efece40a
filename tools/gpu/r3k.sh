set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out/r3k
O=gpurun_out/r3k/cfg2_k_sweep.txt
for K in 5 4 6 7 8 5; do
  echo "== K=$K" >> $O; MVOLPS_LZ_K=$K timeout 120 python profiles/prof_solve.py --iters 100000 --reps 3 2>&1 | tail -1 >> $O
done
for G in 148 144 128 120 74; do
  echo "== G=$G (K=5)" >> $O; timeout 120 python profiles/prof_solve.py --iters 100000 --reps 3 --G $G 2>&1 | tail -1 >> $O
done
cat $O
