export B200HEAD_PDL=0 B200HEAD_ASTAT=0
for v in stamp x_notma x_noepi x_both; do
  B200HEAD_LIB=$PWD/profiles/microbench/libb200head_$v.so timeout 120 python profiles/microbench/epi_timeline.py >> gpurun_out/timeline3.txt 2>&1
done
