# round 2, one-GPU record: full GPU suite, smoke, default bench, launch list, ncu --set full of the hot kernels, soft-DTW phase cycles
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -12 > gpurun_out/r2f_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r2f_smoke.log 2>&1; echo "rc=$?" >> gpurun_out/r2f_smoke.log
timeout 600 python bench.py > gpurun_out/r2f_bench.json 2> gpurun_out/r2f_bench.err; echo "rc=$?" >> gpurun_out/r2f_bench.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2f_launches.csv \
    python bench.py --no-cpu --no-sustained --steps 2 --warmup 3 > gpurun_out/r2f_launches.log 2>&1
timeout 600 ncu --set full --import-source on --clock-control none -k regex:'gemm_tn_2cta' -s 8 -c 4 -o gpurun_out/r2_full_milnce -f \
    python bench.py --no-cpu --no-sustained --no-sdtw --no-graph --steps 2 --warmup 3 > gpurun_out/r2f_full_milnce.log 2>&1
timeout 600 ncu --set full --import-source on --clock-control none -k regex:'fwd_kernel|bwd_kernel' -s 6 -c 2 -o gpurun_out/r2_full_sdtw -f \
    python profiles/sdtw_prof.py > gpurun_out/r2f_full_sdtw.log 2>&1
B200HEAD_LIB=profiles/microbench/libb200head_prof.so timeout 120 python profiles/microbench/sdtw_phases.py > gpurun_out/r2_sdtw_phase_cycles.txt 2>&1
