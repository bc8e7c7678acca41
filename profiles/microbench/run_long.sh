export B200HEAD_PDL=0
nvidia-smi -q -d POWER | grep -i -A8 "Power Readings" | head -12 > gpurun_out/power.txt
for st in 50 500 4000; do python bench.py --no-sdtw --no-cpu --steps $st --warmup 10 > gpurun_out/l_bench_$st.json 2>/dev/null; done
python bench.py --no-sdtw --no-cpu --steps 50 --warmup 10 > gpurun_out/l_bench_50b.json 2>/dev/null
