for mode in "A" "B MSGAT_L2_OLD_APPLY=1" "C MSGAT_L2_DISABLE=1"; do
  set -- $mode
  name=$1; envs=$2
  fails=0
  for i in 1 2 3 4 5 6; do
    env $envs python -m pytest tests/test_gpu_parity.py -q -x -k "pemsd7_full_batch" > /tmp/o.txt 2>&1 || { fails=$((fails+1)); grep -E "^E  " /tmp/o.txt | head -2; }
  done
  echo "mode $name: $fails / 6 failed"
done
