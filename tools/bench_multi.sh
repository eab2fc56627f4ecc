# developer helper: bench.py on N GPUs of one box the way the driver launches it (run through gpurun --gpus N)
N=$1
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/final_bench_${N}gpu.json 2> gpurun_out/final_bench_${N}gpu.err
tail -c 300 gpurun_out/final_bench_${N}gpu.err
head -c 700 gpurun_out/final_bench_${N}gpu.json; echo
