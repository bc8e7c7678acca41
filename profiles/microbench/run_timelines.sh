python -m pytest tests/test_gpu_gemm_core.py tests/test_gpu_milnce.py -x -q 2>&1 | tail -3 > gpurun_out/q_pytest.log
for v in stamp; do
  B200HEAD_LIB=$PWD/profiles/microbench/libb200head_$v.so python profiles/microbench/epi_timeline.py >> gpurun_out/timeline2.txt 2>&1
done
B200HEAD_LIB=$PWD/profiles/microbench/libb200head_stamp.so python profiles/microbench/epi_timeline.py -v > gpurun_out/timeline2_v.txt 2>&1
python bench.py --no-sdtw --no-cpu --steps 50 --warmup 10 > gpurun_out/q_bench.json 2> gpurun_out/q_bench.err
