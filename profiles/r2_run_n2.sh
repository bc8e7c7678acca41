# inside gpurun --gpus 2: multi-GPU tests, bench at N=2, config-5 harness at 1 and 2 GPUs
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_gpu_multi.py -x -q 2>&1 | tail -4 > gpurun_out/r2n2_pytest.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 \
    bench.py --gpus 2 --steps 50 --warmup 10 > gpurun_out/r2n2_bench.json 2> gpurun_out/r2n2_bench.err
echo "rc=$?" >> gpurun_out/r2n2_bench.err
timeout 400 python harness/config5.py --steps 6 --warmup 2 > gpurun_out/r2_config5_1gpu.json 2> gpurun_out/r2_config5_1gpu.err
echo "rc=$?" >> gpurun_out/r2_config5_1gpu.err
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29544 \
    harness/config5.py --steps 6 --warmup 2 > gpurun_out/r2_config5_2gpu.json 2> gpurun_out/r2_config5_2gpu.err
echo "rc=$?" >> gpurun_out/r2_config5_2gpu.err
