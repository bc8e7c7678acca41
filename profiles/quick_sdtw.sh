python -m pytest tests/test_gpu_sdtw.py tests/test_zz_no_alloc.py -x -q 2>&1 | tail -6 > gpurun_out/s_pytest.log
python bench.py --no-cpu --steps 20 --warmup 5 > gpurun_out/s_bench.json 2> gpurun_out/s_bench.err
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'fwd_kernel|bwd_kernel' -c 12 --csv --log-file gpurun_out/s_launches.csv \
    python bench.py --no-cpu --steps 2 --warmup 1 > gpurun_out/s_ncu.log 2>&1
