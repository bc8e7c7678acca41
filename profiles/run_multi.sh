# usage: run_multi.sh N   (inside gpurun --gpus N)
N=$1
python -m pytest tests/test_gpu_multi.py tests/test_gpu_milnce.py -x -q 2>&1 | tail -4 > gpurun_out/m_pytest_$N.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 \
    bench.py --gpus $N --steps 50 --warmup 10 > gpurun_out/m_bench_$N.json 2> gpurun_out/m_bench_$N.err
echo "rc=$?" >> gpurun_out/m_bench_$N.err
