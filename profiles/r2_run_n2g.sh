mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
A="--gpus 2 --steps 50 --warmup 10 --no-sdtw --no-sustained"
timeout 150 $TR --master-port 29535 profiles/ab_slot_barrier.py without $A > gpurun_out/r2n2g_without.json 2> gpurun_out/r2n2g_without.err
timeout 150 $TR --master-port 29536 profiles/ab_slot_barrier.py with $A > gpurun_out/r2n2g_with.json 2> gpurun_out/r2n2g_with.err
timeout 150 $TR --master-port 29537 profiles/ab_slot_barrier.py without $A > gpurun_out/r2n2g_without2.json 2> gpurun_out/r2n2g_without2.err
