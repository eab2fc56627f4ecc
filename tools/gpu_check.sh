# developer helper: GPU parity tests + smoke + one bench run with the per-kernel table (run through gpurun)
set -o pipefail
mkdir -p gpurun_out
TAG=${1:-check}
rm -f gpurun_out/parity_notes.log; timeout 600 python -m pytest tests -m gpu -q --timeout 90 --timeout-method thread 2>&1 | tail -40 > gpurun_out/${TAG}_pytest.log
cat gpurun_out/${TAG}_pytest.log; cat gpurun_out/parity_notes.log 2>/dev/null | cut -c1-400
timeout 200 python __graft_entry__.py smoke > gpurun_out/${TAG}_smoke.log 2>&1; tail -3 gpurun_out/${TAG}_smoke.log
timeout 300 python bench.py --no-cpu-baseline > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err || tail -5 gpurun_out/${TAG}_bench.err
python - <<PY
import json
try:
    d=json.load(open("gpurun_out/${TAG}_bench.json"))
    print(d["value"], d["ms_per_step"], d["e2e"]["value"], d["roofline"]["my_kernels_ms_per_step"], d["gpu_launches"])
    for k,v in d["kernels"].items():
        print(f'{k:32s} {v["launches_per_step"]:5.1f} x {v["ms_per_launch"]*1e3:8.1f} us  {v["share"]*100:5.1f}%  {v["GBps"]}')
except Exception as e:
    print("bench parse failed", e)
PY
