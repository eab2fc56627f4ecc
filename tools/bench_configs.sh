# developer helper: bench lines of the other BASELINE.json configurations (1 GPU)
mkdir -p gpurun_out
for cfg in pemsd8 pemsd4 pemsd3 n4096; do
  timeout 300 python bench.py --config $cfg --no-cpu-baseline --steps 10 --warmup 3 > gpurun_out/cfg_$cfg.json 2> gpurun_out/cfg_$cfg.err
  python - <<PY
import json
d=json.load(open("gpurun_out/cfg_$cfg.json"))
print("$cfg", round(d["value"],1), round(d["ms_per_step"],2), d["roofline"]["kernel"], round(d["roofline"]["frac"],3))
ks=sorted(d["kernels"].items(), key=lambda kv:-kv[1]["share"])[:6]
for k,v in ks: print("   ", k, round(v["ms_per_launch"]*1e3,1), round(v["share"]*100,1))
PY
done
