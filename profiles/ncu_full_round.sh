# one ncu --set full capture of the six hot kernels of a default bench step (after a plain run exited 0)
python bench.py --no-cpu --steps 2 --warmup 1 > gpurun_out/f_plain.json 2> gpurun_out/f_plain.err || exit 1
ncu --set full --import-source on --clock-control none -k regex:'gemm_tn_2cta|fwd_kernel|bwd_kernel' -s 6 -c 6 \
    -o gpurun_out/r1_full_hot -f python bench.py --no-cpu --steps 2 --warmup 1 > gpurun_out/f_ncu.log 2>&1
ls -la gpurun_out/r1_full_hot.ncu-rep
