# inside gpurun --gpus 8: timeline of the default path and bench at N=8, then N=4 on four of the GPUs
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 300 $TR --nproc-per-node 8 --master-port 29511 profiles/multi_timeline.py --only pull --out gpurun_out/r2_multi_timeline_8.txt > gpurun_out/r2n8_timeline_8.log 2>&1
timeout 300 $TR --nproc-per-node 8 --master-port 29533 bench.py --gpus 8 --steps 50 --warmup 10 --no-sustained > gpurun_out/r2n8_bench_8.json 2> gpurun_out/r2n8_bench_8.err
timeout 200 $TR --nproc-per-node 4 --master-port 29534 bench.py --gpus 4 --steps 50 --warmup 10 --no-sdtw --no-sustained > gpurun_out/r2n8_bench_4.json 2> gpurun_out/r2n8_bench_4.err
timeout 200 $TR --nproc-per-node 4 --master-port 29512 profiles/multi_timeline.py --only pull --steps 15 --out gpurun_out/r2_multi_timeline_4.txt > gpurun_out/r2n8_timeline_4.log 2>&1
