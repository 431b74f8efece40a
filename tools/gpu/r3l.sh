set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out/r3l
O=gpurun_out/r3l/ring_pf.txt
echo "== product (K=12, registers + L2 prefetch)" >> $O; timeout 120 python profiles/prof_batch.py 592 256 512 >> $O 2>&1
for v in rpf2 rpf4; do
for K in 10 8; do
  echo "== $v ring K=$K" >> $O; MVOLPS_LIB=build/variants/$v.so MVOLPS_LZ_K=$K timeout 120 python profiles/prof_batch.py 592 256 512 >> $O 2>&1
done
done
echo "== product ring K=10 (no prefetch)" >> $O; MVOLPS_LZ_K=10 timeout 120 python profiles/prof_batch.py 592 256 512 >> $O 2>&1
grep "==\|LPs" $O | cut -c1-150
