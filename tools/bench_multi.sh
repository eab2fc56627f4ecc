# developer helper: bench.py on N GPUs of one box the way the driver launches it (run through gpurun --gpus N): weak scaling
# (with the strong sub-record) and the exit codes of the ranks
N=$1; TAG=${2:-r02}
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/${TAG}_bench_${N}gpu.json 2> gpurun_out/${TAG}_bench_${N}gpu.err
echo "exit code $?" | tee -a gpurun_out/${TAG}_bench_${N}gpu.err
tail -c 400 gpurun_out/${TAG}_bench_${N}gpu.err
python - <<PY
import json
d=json.loads([l for l in open("gpurun_out/${TAG}_bench_${N}gpu.json") if l.startswith("{")][-1])
print({k: d[k] for k in ("value","n_gpus","ms_per_step","scaling","gpu_launches")}, d["e2e"]["value"], d.get("strong"))
PY
