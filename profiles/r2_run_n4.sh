# inside gpurun --gpus 4: finer forward timeline of the peer path at W=4 and W=2
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 200 $TR --nproc-per-node 4 --master-port 29512 profiles/multi_timeline.py --only "pull" --out gpurun_out/r2_multi_timeline_4_fine.txt > gpurun_out/r2n4_timeline.log 2>&1
echo "rc=$?" >> gpurun_out/r2n4_timeline.log
timeout 200 $TR --nproc-per-node 2 --master-port 29513 profiles/multi_timeline.py --only "pull" --out gpurun_out/r2_multi_timeline_2_fine.txt > gpurun_out/r2n4_timeline2.log 2>&1
