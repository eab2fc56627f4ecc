# Round-end measurement on one B200 (run through gpurun): parity tests, smoke, bench (both arms), the other
# BASELINE.json configurations, the ncu launch list of the bench command and ncu --set full captures of the
# top kernels (each after the same command has exited 0 without ncu).  Everything lands in gpurun_out/final_*.
set -o pipefail
mkdir -p gpurun_out
timeout 400 python -m pytest tests -m gpu -q 2>&1 | tail -4 > gpurun_out/final_pytest_gpu.log
timeout 120 python __graft_entry__.py smoke > gpurun_out/final_smoke.log 2>&1
timeout 300 python bench.py > gpurun_out/final_bench.json 2> gpurun_out/final_bench.err
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/final_bench_reference_arm.json 2> gpurun_out/final_ref.err
timeout 200 python bench.py --precision tf32 --no-cpu-baseline > gpurun_out/final_bench_tf32.json 2> gpurun_out/final_tf32.err
for cfg in pemsd8 pemsd4 pemsd3 n4096; do
  timeout 300 python bench.py --config $cfg --no-cpu-baseline --steps 10 --warmup 3 > gpurun_out/final_bench_cfg_$cfg.json 2> gpurun_out/final_cfg_$cfg.err
done
BENCH="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --graph off"
timeout 200 $BENCH > gpurun_out/final_plain.json 2> gpurun_out/final_plain.err || exit 1
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/final_ncu_launches.csv $BENCH > gpurun_out/final_ncu_launches.log 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:"k_tc_dgrad|k_tc_wgrad|k_tc_apply3|k_bwd_main" --launch-skip 4 -c 5 -f -o gpurun_out/final_ncu_a $BENCH > gpurun_out/final_ncu_a.log 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:"k_head_fwd12|k_sstats|k_prep|k_apply" -c 6 -f -o gpurun_out/final_ncu_b $BENCH > gpurun_out/final_ncu_b.log 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:"k_head_bwd12|k_scol|k_bwd_fix" -c 5 -f -o gpurun_out/final_ncu_c $BENCH > gpurun_out/final_ncu_c.log 2>&1
cat gpurun_out/final_pytest_gpu.log gpurun_out/final_smoke.log
head -c 400 gpurun_out/final_bench.json; echo
ls -la gpurun_out/final_*
