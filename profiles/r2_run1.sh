# round 2, first GPU run: full GPU test suite, smoke, default bench, reference arm, host-overhead probe
set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -25 > gpurun_out/r2a_pytest.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r2a_smoke.log 2>&1; echo "rc=$?" >> gpurun_out/r2a_smoke.log
python bench.py > gpurun_out/r2a_bench.json 2> gpurun_out/r2a_bench.err; echo "rc=$?" >> gpurun_out/r2a_bench.err
python bench.py --impl reference > gpurun_out/r2a_ref.json 2> gpurun_out/r2a_ref.err; echo "rc=$?" >> gpurun_out/r2a_ref.err
python profiles/host_overhead.py > gpurun_out/r2a_host.log 2>&1
nproc >> gpurun_out/r2a_host.log; free -g >> gpurun_out/r2a_host.log
