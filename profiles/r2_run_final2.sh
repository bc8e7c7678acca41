# round 2, closing one-GPU check: full GPU suite, smoke, default bench
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -12 > gpurun_out/r2g_pytest.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r2g_smoke.log 2>&1; echo "rc=$?" >> gpurun_out/r2g_smoke.log
timeout 400 python bench.py > gpurun_out/r2g_bench.json 2> gpurun_out/r2g_bench.err; echo "rc=$?" >> gpurun_out/r2g_bench.err
