# inside gpurun --gpus 2: peer-path tests (single-GPU emulation and 2-GPU), peer timelines, bench at N=2, config-5 parity
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_peer_single.py tests/test_gpu_multi.py -x -q 2>&1 | tail -8 > gpurun_out/r2n2b_pytest.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
    profiles/multi_timeline.py --only peer --out gpurun_out/r2_multi_timeline_2_peer.txt > gpurun_out/r2n2b_timeline.log 2>&1
echo "rc=$?" >> gpurun_out/r2n2b_timeline.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 \
    bench.py --gpus 2 --steps 50 --warmup 10 --no-sdtw --no-sustained > gpurun_out/r2n2b_bench.json 2> gpurun_out/r2n2b_bench.err
echo "rc=$?" >> gpurun_out/r2n2b_bench.err
timeout 300 python harness/config5.py --steps 4 --warmup 2 > gpurun_out/r2_config5_1gpu.json 2> gpurun_out/r2_config5_1gpu.err
echo "rc=$?" >> gpurun_out/r2_config5_1gpu.err
