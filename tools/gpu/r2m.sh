set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out/r2m
O=gpurun_out/r2m/prof_batch.txt
MVOLPS_LIB=build/variants/prof.so timeout 120 python profiles/prof_batch.py 592 256 512 >> $O 2>&1
echo "---- 296 LPs (one per CTA, no second wave)" >> $O
MVOLPS_LIB=build/variants/prof.so timeout 120 python profiles/prof_batch.py 296 256 512 >> $O 2>&1
echo "---- 148 LPs" >> $O
MVOLPS_LIB=build/variants/prof.so timeout 120 python profiles/prof_batch.py 148 256 512 >> $O 2>&1
cat $O
