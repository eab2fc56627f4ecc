# developer helper: GPU parity tests + one bench run with the per-kernel table (run through gpurun)
set -o pipefail
mkdir -p gpurun_out
timeout 300 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 200 python bench.py --no-cpu-baseline > gpurun_out/check_bench.json 2> gpurun_out/check_bench.err
python - <<PY
import json
d=json.load(open("gpurun_out/check_bench.json"))
print(d["value"], d["ms_per_step"], d["e2e"]["value"], d["roofline"]["my_kernels_ms_per_step"])
for k,v in d["kernels"].items():
    print(f'{k:32s} {v["ms_per_launch"]*1e3:8.1f} us  {v["share"]*100:5.1f}%  {v["GBps"]}')
PY
