# usage: run_multi_bench.sh N   (inside gpurun --gpus N): the driver's launch line for N ranks
N=$1
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 \
    bench.py --gpus $N --steps 50 --warmup 10 > gpurun_out/m_bench_$N.json 2> gpurun_out/m_bench_$N.err
echo "rc=$?" >> gpurun_out/m_bench_$N.err
