# Round-end measurement on one B200 (run through gpurun): parity tests, smoke, bench (both arms), TF32 mode, the other
# BASELINE.json configurations, and the ncu launch list of the bench command (after the same command has exited 0
# without ncu).  Everything lands in gpurun_out/final_*; ncu --set full captures are taken one per gpurun call with
# tools/ncu_one.sh.
set -o pipefail
mkdir -p gpurun_out
rm -f gpurun_out/parity_notes.log
timeout 600 python -m pytest tests -m gpu -q --timeout 120 --timeout-method thread 2>&1 | tail -12 > gpurun_out/final_pytest_gpu.log
cp gpurun_out/parity_notes.log gpurun_out/final_parity_notes.log 2>/dev/null
timeout 120 python __graft_entry__.py smoke > gpurun_out/final_smoke.log 2>&1
timeout 300 python bench.py > gpurun_out/final_bench.json 2> gpurun_out/final_bench.err
timeout 400 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/final_bench_reference_arm.json 2> gpurun_out/final_ref.err
timeout 200 python bench.py --precision tf32 --no-cpu-baseline > gpurun_out/final_bench_tf32.json 2> gpurun_out/final_tf32.err
for cfg in pemsd8 pemsd4 pemsd3 n4096; do
  timeout 300 python bench.py --config $cfg --no-cpu-baseline --steps 10 --warmup 3 > gpurun_out/final_bench_cfg_$cfg.json 2> gpurun_out/final_cfg_$cfg.err
done
BENCH="python bench.py --steps 2 --warmup 3 --windows 1 --no-cpu-baseline --graph off --profile-steps 1"
timeout 200 $BENCH > gpurun_out/final_plain.json 2> gpurun_out/final_plain.err || exit 1
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/final_ncu_launches.csv $BENCH > gpurun_out/final_ncu_launches.log 2>&1
cat gpurun_out/final_pytest_gpu.log | tail -3; cat gpurun_out/final_smoke.log | tail -1
for f in final_bench final_bench_tf32 final_bench_cfg_pemsd8 final_bench_cfg_pemsd4 final_bench_cfg_pemsd3 final_bench_cfg_n4096 final_bench_reference_arm; do
  python - <<PY
import json
try:
    d=json.loads([l for l in open("gpurun_out/$f.json") if l.startswith("{")][-1])
    print("$f", round(d["value"],1), round(d["ms_per_step"],3), d.get("gpu_launches"), (d.get("cpu_baseline") or {}).get("kind"))
except Exception as e:
    print("$f FAILED", e)
PY
done
ls -la gpurun_out/final_ncu_launches.csv
