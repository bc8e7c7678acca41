mkdir -p gpurun_out
timeout 200 python -m pytest tests/test_gpu_multi.py -x -q 2>&1 | tail -3 > gpurun_out/r2n2f_pytest.log
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29535 bench.py --gpus 2 --steps 200 --warmup 20 --no-sdtw --no-sustained > gpurun_out/r2n2f_bench.json 2> gpurun_out/r2n2f_bench.err
