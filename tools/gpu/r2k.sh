set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out/r2k
O=gpurun_out/r2k/ksweep.txt
echo "product K=8" >> $O; timeout 120 python profiles/prof_batch.py 592 256 512 >> $O 2>&1
for lib in k14u4 k14u2; do
  for K in 8 10 12 14; do
    echo "$lib K=$K" >> $O
    MVOLPS_LIB=build/variants/$lib.so MVOLPS_LZ_K=$K timeout 120 python profiles/prof_batch.py 592 256 512 >> $O 2>&1
  done
done
echo "k14u4 K=14 4096 LPs" >> $O; MVOLPS_LIB=build/variants/k14u4.so MVOLPS_LZ_K=14 timeout 200 python profiles/prof_batch.py 4096 256 512 >> $O 2>&1
echo "product K=8 4096 LPs" >> $O; timeout 200 python profiles/prof_batch.py 4096 256 512 >> $O 2>&1
cat $O
