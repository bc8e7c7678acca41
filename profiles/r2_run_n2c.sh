# inside gpurun --gpus 2: MIL-NCE tests (single-GPU, incl. rank emulation), 2-GPU tests, peer timelines, bench at N=2
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_multi.py -x -q 2>&1 | tail -8 > gpurun_out/r2n2c_pytest.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
    profiles/multi_timeline.py --only peer --out gpurun_out/r2_multi_timeline_2_peer.txt > gpurun_out/r2n2c_timeline.log 2>&1
echo "rc=$?" >> gpurun_out/r2n2c_timeline.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 \
    bench.py --gpus 2 --steps 50 --warmup 10 --no-sdtw --no-sustained > gpurun_out/r2n2c_bench.json 2> gpurun_out/r2n2c_bench.err
echo "rc=$?" >> gpurun_out/r2n2c_bench.err
