# inside gpurun --gpus 8: per-phase timeline of the default path (W=8, W=4), bench at N=8 and N=4, config-5 at 8 x 32 = 256
# (all variants at W=8 were recorded earlier in the round: profiles/r2_multi_timeline_8_all_variants.txt)
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 300 $TR --nproc-per-node 8 --master-port 29511 profiles/multi_timeline.py --only pull --out gpurun_out/r2_multi_timeline_8.txt > gpurun_out/r2n8_timeline_8.log 2>&1
echo "rc=$?" >> gpurun_out/r2n8_timeline_8.log
timeout 300 $TR --nproc-per-node 8 --master-port 29533 bench.py --gpus 8 --steps 50 --warmup 10 > gpurun_out/r2n8_bench_8.json 2> gpurun_out/r2n8_bench_8.err
echo "rc=$?" >> gpurun_out/r2n8_bench_8.err
timeout 300 $TR --nproc-per-node 8 --master-port 29544 harness/config5.py --steps 6 --warmup 2 > gpurun_out/r2_config5_8gpu.json 2> gpurun_out/r2_config5_8gpu.err
echo "rc=$?" >> gpurun_out/r2_config5_8gpu.err
timeout 200 $TR --nproc-per-node 4 --master-port 29512 profiles/multi_timeline.py --only pull --out gpurun_out/r2_multi_timeline_4.txt > gpurun_out/r2n8_timeline_4.log 2>&1
echo "rc=$?" >> gpurun_out/r2n8_timeline_4.log
timeout 200 $TR --nproc-per-node 4 --master-port 29534 bench.py --gpus 4 --steps 50 --warmup 10 --no-sdtw --no-sustained > gpurun_out/r2n8_bench_4.json 2> gpurun_out/r2n8_bench_4.err
echo "rc=$?" >> gpurun_out/r2n8_bench_4.err
timeout 200 $TR --nproc-per-node 2 --master-port 29535 bench.py --gpus 2 --steps 50 --warmup 10 --no-sdtw --no-sustained > gpurun_out/r2n8_bench_2.json 2> gpurun_out/r2n8_bench_2.err
echo "rc=$?" >> gpurun_out/r2n8_bench_2.err
