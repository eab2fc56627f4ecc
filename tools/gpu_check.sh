set -o pipefail
timeout 300 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 200 python bench.py --no-cpu-baseline > gpurun_out/s19_bench3.json 2> gpurun_out/s19_bench3.err
python - <<PY
import json
d=json.load(open("gpurun_out/s19_bench3.json"))
print(d["value"], d["ms_per_step"], d["e2e"]["value"], d["roofline"]["my_kernels_ms_per_step"])
for k,v in d["kernels"].items():
    if "head" in k or "reduce@C72O12" in k: print(k, v)
PY
