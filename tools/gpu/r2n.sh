set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out/r2n
(time timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "streaming") > gpurun_out/r2n/pytest_stream.log 2>&1; echo "stream rc=$?"; tail -n 15 gpurun_out/r2n/pytest_stream.log
O=gpurun_out/r2n/ring_sweep.txt
echo "product ring" >> $O; timeout 120 python profiles/prof_batch.py 592 256 512 >> $O 2>&1
echo "product no ring" >> $O; MVOLPS_LZ_RING=0 timeout 120 python profiles/prof_batch.py 592 256 512 >> $O 2>&1
for K in 10 12 14; do
  echo "k14u4 ring K=$K" >> $O
  MVOLPS_LIB=build/variants/k14u4.so MVOLPS_LZ_K=$K timeout 120 python profiles/prof_batch.py 592 256 512 >> $O 2>&1
done
echo "product ring 4096" >> $O; timeout 200 python profiles/prof_batch.py 4096 256 512 >> $O 2>&1
echo "prof ring" >> $O; MVOLPS_LIB=build/variants/prof.so timeout 120 python profiles/prof_batch.py 592 256 512 >> $O 2>&1
echo "product ring 148 LPs" >> $O; timeout 120 python profiles/prof_batch.py 148 256 512 >> $O 2>&1
echo "product no ring 148 LPs" >> $O; MVOLPS_LZ_RING=0 timeout 120 python profiles/prof_batch.py 148 256 512 >> $O 2>&1
cat $O
