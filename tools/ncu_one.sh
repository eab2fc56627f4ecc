# developer helper: ncu --set full capture of ONE kernel of the bench step (run through gpurun): [BENCH_EXTRA='--precision tf32'] tools/ncu_one.sh <regex> <tag> [skip]
set -o pipefail
mkdir -p gpurun_out
K=$1; TAG=$2; SKIP=${3:-0}
BENCH="python bench.py --steps 1 --warmup 3 --windows 1 --no-cpu-baseline --graph off --profile-steps 1 ${BENCH_EXTRA:-}"
timeout 300 $BENCH > gpurun_out/${TAG}_plain.json 2> gpurun_out/${TAG}_plain.err || { tail -5 gpurun_out/${TAG}_plain.err; exit 1; }
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"$K" --launch-skip $SKIP -c 1 -f -o gpurun_out/${TAG} $BENCH > gpurun_out/${TAG}_ncu.log 2>&1
tail -3 gpurun_out/${TAG}_ncu.log
ls -la gpurun_out/${TAG}.ncu-rep
