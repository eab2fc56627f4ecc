mkdir -p gpurun_out
timeout 300 ncu --set full --clock-control none --import-source on -k regex:"k_head_(fwd|bwd)12" --launch-skip 4 -c 2 -o gpurun_out/s19_head -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline --graph off > gpurun_out/s19_ncu_head.log 2>&1
tail -3 gpurun_out/s19_ncu_head.log
