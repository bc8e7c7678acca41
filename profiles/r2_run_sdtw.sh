# soft-DTW kernels: parity tests, then the bench line's soft-DTW leg and a launch list
python -m pytest tests/test_gpu_sdtw.py tests/test_gpu_guards.py tests/test_zz_no_alloc.py -x -q 2>&1 | tail -15 > gpurun_out/r2s_pytest.log
python bench.py --no-cpu --no-sustained --steps 5 --warmup 3 > gpurun_out/r2s_bench.json 2> gpurun_out/r2s_bench.err
python - <<'PY' > gpurun_out/r2s_sdtw.txt 2>&1
import json
d = json.load(open("gpurun_out/r2s_bench.json"))
print(json.dumps(d["sdtw"], indent=1))
PY
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"fwd_kernel|bwd_kernel" -c 12 --csv --log-file gpurun_out/r2s_launches.csv \
    python profiles/sdtw_prof.py > gpurun_out/r2s_ncu.log 2>&1
