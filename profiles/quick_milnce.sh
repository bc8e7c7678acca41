# quick check of the MIL-NCE path: parity tests, a bench line, then the per-kernel launch list
python -m pytest tests/test_gpu_gemm_core.py tests/test_gpu_milnce.py tests/test_gpu_guards.py -x -q 2>&1 | tail -8 > gpurun_out/q_pytest.log
python bench.py --no-sdtw --no-cpu --steps 50 --warmup 10 > gpurun_out/q_bench.json 2> gpurun_out/q_bench.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 30 --csv --log-file gpurun_out/q_launches.csv \
    python bench.py --no-sdtw --no-cpu --steps 2 --warmup 1 > gpurun_out/q_ncu.log 2>&1
