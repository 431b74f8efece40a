set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out/r2p
(time timeout 1500 python -m pytest tests -m gpu -x -q) > gpurun_out/r2p/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 6 gpurun_out/r2p/pytest_gpu.log
O=gpurun_out/r2p/ring_sweep.txt
echo "product ring" >> $O; timeout 120 python profiles/prof_batch.py 592 256 512 >> $O 2>&1
echo "product no ring" >> $O; MVOLPS_LZ_RING=0 timeout 120 python profiles/prof_batch.py 592 256 512 >> $O 2>&1
for K in 10 12 14; do
  echo "k14u4 ring K=$K" >> $O
  MVOLPS_LIB=build/variants/k14u4.so MVOLPS_LZ_K=$K timeout 120 python profiles/prof_batch.py 592 256 512 >> $O 2>&1
done
echo "k14u4 no ring K=12" >> $O; MVOLPS_LZ_RING=0 MVOLPS_LIB=build/variants/k14u4.so MVOLPS_LZ_K=12 timeout 120 python profiles/prof_batch.py 592 256 512 >> $O 2>&1
echo "product ring 4096" >> $O; timeout 200 python profiles/prof_batch.py 4096 256 512 >> $O 2>&1
echo "prof ring" >> $O; MVOLPS_LIB=build/variants/prof.so timeout 120 python profiles/prof_batch.py 592 256 512 >> $O 2>&1
echo "product ring 148 LPs" >> $O; timeout 120 python profiles/prof_batch.py 148 256 512 >> $O 2>&1
cat $O | grep -v "^ "
(time timeout 300 python bench.py --headline-only --steps 5 --warmup 3) > gpurun_out/r2p/bench_headline.log 2>&1; tail -c 600 gpurun_out/r2p/bench_headline.log
