set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > gpurun_out/v_pytest.log
python bench.py > gpurun_out/v_bench.json 2> gpurun_out/v_bench.err; echo "rc=$?" >> gpurun_out/v_bench.err
wc -l gpurun_out/v_bench.json
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/v_ref.json 2> gpurun_out/v_ref.err; echo "rc=$?" >> gpurun_out/v_ref.err
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/v_smoke.log 2>&1; echo "rc=$?" >> gpurun_out/v_smoke.log
