set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out/r2t
O=gpurun_out/r2t/prof_pf4.txt
MVOLPS_LIB=build/variants/prof.so timeout 120 python profiles/prof_batch.py 592 256 512 >> $O 2>&1
echo "---- K=8 ring" >> $O
MVOLPS_LZ_K=8 MVOLPS_LIB=build/variants/prof.so timeout 120 python profiles/prof_batch.py 592 256 512 >> $O 2>&1
cat $O
timeout 300 ncu --set full --clock-control none --import-source on -k regex:simplex_ps_stream -c 1 -o gpurun_out/r2t/stream_kernel python profiles/prof_batch.py 592 256 512 > gpurun_out/r2t/ncu.log 2>&1; tail -3 gpurun_out/r2t/ncu.log
