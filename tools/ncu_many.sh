# developer helper: ncu --set full captures of several kernels of the bench step in ONE pass each (run through gpurun):
#   tools/ncu_many.sh <tag> <regex1> [<regex2> ...]     -> gpurun_out/<tag>_<i>.ncu-rep
set -o pipefail
mkdir -p gpurun_out
TAG=$1; shift
BENCH="python bench.py --steps 1 --warmup 3 --windows 1 --no-cpu-baseline --graph off --profile-steps 1 ${BENCH_EXTRA:-}"
i=0
for K in "$@"; do
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:"$K" --launch-skip ${SKIP:-0} -c ${COUNT:-1} -f -o gpurun_out/${TAG}_$i $BENCH > gpurun_out/${TAG}_${i}_ncu.log 2>&1
  tail -1 gpurun_out/${TAG}_${i}_ncu.log
  i=$((i+1))
done
