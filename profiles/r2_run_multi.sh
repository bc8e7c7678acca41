# usage: r2_run_multi.sh N   (inside gpurun --gpus N): NCCL parity test, per-phase timeline, bench
N=$1
python -m pytest tests/test_gpu_multi.py -x -q 2>&1 | tail -4 > gpurun_out/r2m_pytest_$N.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
    profiles/multi_timeline.py --out gpurun_out/r2_multi_timeline_$N.txt > gpurun_out/r2m_timeline_$N.log 2>&1
echo "rc=$?" >> gpurun_out/r2m_timeline_$N.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 \
    bench.py --gpus $N --steps 50 --warmup 10 > gpurun_out/r2m_bench_$N.json 2> gpurun_out/r2m_bench_$N.err
echo "rc=$?" >> gpurun_out/r2m_bench_$N.err
