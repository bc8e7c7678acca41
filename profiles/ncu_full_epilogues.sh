set -x
B200HEAD_PDL=0 python bench.py --no-sdtw --no-cpu --steps 2 --warmup 1 > gpurun_out/f_plain.json 2> gpurun_out/f_plain.err || exit 1
B200HEAD_PDL=0 ncu --set full --import-source on --clock-control none -k regex:gemm_tn_2cta -c 2 -o gpurun_out/r1_full_epi -f \
    python bench.py --no-sdtw --no-cpu --steps 2 --warmup 1 > gpurun_out/f_ncu.log 2>&1
ls -la gpurun_out/*.ncu-rep
