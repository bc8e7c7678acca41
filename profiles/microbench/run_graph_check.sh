python -m pytest tests/test_gpu_milnce.py -x -q 2>&1 | tail -4 > gpurun_out/q_pytest.log
python bench.py --no-sdtw --no-cpu --steps 50 --warmup 10 > gpurun_out/q_bench.json 2> gpurun_out/q_bench.err
python bench.py --no-sdtw --no-cpu --steps 50 --warmup 10 --no-graph > gpurun_out/q_bench_nograph.json 2> gpurun_out/q_bench_nograph.err
