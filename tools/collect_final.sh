# developer helper: copy the round-end measurement of gpurun_out/ into profiles/ (tracked) under the r02_final_ prefix
cd "$(dirname "$0")/.."
for f in gpurun_out/final_*.json gpurun_out/final_pytest_gpu.log gpurun_out/final_smoke.log gpurun_out/final_parity_notes.log gpurun_out/final_ncu_launches.csv; do
  [ -s "$f" ] || continue
  b=$(basename "$f"); b=${b#final_}
  case "$f" in
    *.json) grep "^{" "$f" | tail -1 > profiles/r02_final_$b ;;
    *) cp "$f" profiles/r02_final_$b ;;
  esac
done
rm -f profiles/r02_final_plain.json
ls profiles/r02_final_* | wc -l
