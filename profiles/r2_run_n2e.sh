mkdir -p gpurun_out
timeout 200 python -m pytest tests/test_gpu_peer_single.py tests/test_gpu_multi.py -x -q 2>&1 | tail -3 > gpurun_out/r2n2e_pytest.log
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 profiles/multi_timeline.py --only "pull" --out gpurun_out/r2_multi_timeline_2.txt > gpurun_out/r2n2e_timeline.log 2>&1
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29535 bench.py --gpus 2 --steps 50 --warmup 10 --no-sdtw --no-sustained > gpurun_out/r2n8_bench_2.json 2> gpurun_out/r2n8_bench_2.err
