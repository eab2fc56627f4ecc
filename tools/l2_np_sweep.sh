# developer helper: producer-warp sweep of the pipelined layer-2 kernels (run through gpurun)
mkdir -p gpurun_out
for cfg in "3 1" "3 2" "3 3" "2 1" "1 1"; do
  set -- $cfg
  MSGAT_L2_NP_MIX=$1 MSGAT_L2_NP_OUT=$2 timeout 120 python bench.py --steps 10 --warmup 3 --windows 2 --no-cpu-baseline > gpurun_out/np_$1_$2.json 2>/dev/null
  python - <<PY
import json
d=json.load(open("gpurun_out/np_$1_$2.json"))
k=d["kernels"]
print("mix np=$1 out np=$2", round(d["ms_per_step"],3), {n: round(v["ms_per_launch"]*1e3,1) for n,v in k.items() if "k_l2" in n})
PY
done
