# end-of-round verification: full GPU test suite, default bench, reference arm, smoke, fp32 bench, launch list
set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > gpurun_out/v_pytest.log
python bench.py > gpurun_out/v_bench.json 2> gpurun_out/v_bench.err; echo "rc=$?" >> gpurun_out/v_bench.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/v_ref.json 2> gpurun_out/v_ref.err; echo "rc=$?" >> gpurun_out/v_ref.err
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/v_smoke.log 2>&1; echo "rc=$?" >> gpurun_out/v_smoke.log
python bench.py --dtype fp32 --no-sdtw --no-cpu --steps 10 --warmup 3 > gpurun_out/v_bench_fp32.json 2> gpurun_out/v_bench_fp32.err
python bench.py --no-cpu --steps 3000 --warmup 10 > gpurun_out/v_bench_sustained.json 2> gpurun_out/v_bench_sustained.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file gpurun_out/v_launches.csv \
    python bench.py --no-cpu --steps 2 --warmup 1 > gpurun_out/v_ncu.log 2>&1
