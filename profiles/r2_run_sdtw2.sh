# A/B of the early L2 prefetch in the soft-DTW backward: parity tests, then the kernel pair timed over rotating inputs
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_sdtw.py tests/test_gpu_composite.py tests/test_zz_no_alloc.py -x -q 2>&1 | tail -3 > gpurun_out/r2j_pytest.log
timeout 200 python profiles/sdtw_time_rot.py > gpurun_out/r2j_rot_new.txt 2>&1
B200HEAD_LIB=profiles/microbench/libb200head_VARIANT.so timeout 200 python profiles/sdtw_time_rot.py > gpurun_out/r2j_rot_old.txt 2>&1
