# developer helper: ncu --set full of the kernels matching $1 (regex), 2 launches, into gpurun_out/$2.ncu-rep
mkdir -p gpurun_out
timeout 300 ncu --set full --clock-control none --import-source on -k regex:"$1" --launch-skip ${3:-4} -c ${4:-2} -o gpurun_out/$2 -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline --graph off > gpurun_out/$2.log 2>&1
tail -2 gpurun_out/$2.log | cut -c1-200
