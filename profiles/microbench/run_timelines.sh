python -m pytest tests/test_gpu_gemm_core.py tests/test_gpu_milnce.py -x -q 2>&1 | tail -3 > gpurun_out/q_pytest.log
for a in 1 0; do
  B200HEAD_EPIG_ASTAT=$a B200HEAD_LIB=$PWD/profiles/microbench/libb200head_stamp.so python profiles/microbench/epi_timeline.py >> gpurun_out/timeline2.txt 2>&1
  B200HEAD_EPIG_ASTAT=$a python bench.py --no-sdtw --no-cpu --steps 50 --warmup 10 > gpurun_out/q_bench_ga$a.json 2> gpurun_out/q_bench_ga$a.err
done
B200HEAD_EPIG_ASTAT=1 python -m pytest tests/test_gpu_milnce.py -x -q 2>&1 | tail -3 >> gpurun_out/q_pytest.log
