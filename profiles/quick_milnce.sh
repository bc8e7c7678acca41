# quick check of the MIL-NCE path: parity tests, bench lines (PDL on/off, optional variant builds), per-kernel launch list
set -x
python -m pytest tests/test_gpu_gemm_core.py tests/test_gpu_milnce.py -x -q 2>&1 | tail -5 > gpurun_out/q_pytest.log
B200HEAD_PDL=0 python bench.py --no-sdtw --no-cpu --steps 50 --warmup 10 > gpurun_out/q_bench_nopdl.json 2> gpurun_out/q_bench_nopdl.err || exit 1
python bench.py --no-sdtw --no-cpu --steps 50 --warmup 10 > gpurun_out/q_bench.json 2> gpurun_out/q_bench.err
for v in $VARIANTS; do
  B200HEAD_PDL=0 B200HEAD_LIB=$PWD/profiles/microbench/libb200head_$v.so python bench.py --no-sdtw --no-cpu --steps 50 --warmup 10 > gpurun_out/q_bench_$v.json 2> gpurun_out/q_bench_$v.err
done
B200HEAD_PDL=0 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/q_launches.csv \
    python bench.py --no-sdtw --no-cpu --steps 2 --warmup 1 > gpurun_out/q_ncu.log 2>&1
