# inside gpurun --gpus 8: after the rotated staging order - peer tests, timeline of the default path and bench at N=8 and N=4
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 200 python -m pytest tests/test_gpu_peer_single.py tests/test_gpu_multi.py -x -q 2>&1 | tail -4 > gpurun_out/r2n8b_pytest.log
timeout 300 $TR --nproc-per-node 8 --master-port 29511 profiles/multi_timeline.py --only pull --out gpurun_out/r2_multi_timeline_8.txt > gpurun_out/r2n8_timeline_8.log 2>&1
timeout 300 $TR --nproc-per-node 8 --master-port 29533 bench.py --gpus 8 --steps 50 --warmup 10 --no-sdtw --no-sustained > gpurun_out/r2n8_bench_8.json 2> gpurun_out/r2n8_bench_8.err
timeout 200 $TR --nproc-per-node 4 --master-port 29512 profiles/multi_timeline.py --only pull --out gpurun_out/r2_multi_timeline_4.txt > gpurun_out/r2n8_timeline_4.log 2>&1
timeout 200 $TR --nproc-per-node 4 --master-port 29534 bench.py --gpus 4 --steps 50 --warmup 10 --no-sdtw --no-sustained > gpurun_out/r2n8_bench_4.json 2> gpurun_out/r2n8_bench_4.err
timeout 200 $TR --nproc-per-node 2 --master-port 29535 bench.py --gpus 2 --steps 50 --warmup 10 --no-sdtw --no-sustained > gpurun_out/r2n8_bench_2.json 2> gpurun_out/r2n8_bench_2.err
