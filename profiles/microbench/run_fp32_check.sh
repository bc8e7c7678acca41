python -m pytest tests/test_gpu_gemm_core.py tests/test_gpu_milnce.py -x -q 2>&1 | tail -8 > gpurun_out/q_pytest.log
python bench.py --dtype fp32 --no-sdtw --no-cpu --steps 10 --warmup 3 > gpurun_out/q_bench_fp32.json 2> gpurun_out/q_bench_fp32.err
