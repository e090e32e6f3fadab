#!/bin/bash
# A/B of plan toggles (environment variables) on the same box: bash tools/ab_env.sh "" "AY_RES_KB=128" ...
O=gpurun_out/ab; mkdir -p $O
i=0
for v in "$@"; do
  i=$((i+1))
  env $v timeout 150 python tools/layer_times.py --batch 256 --reps 5 --out $O/lt_env$i.txt > $O/lt_env$i.log 2>&1
  echo "[$v] $(grep 'sum us' $O/lt_env$i.txt) | $(grep 'whole forward' $O/lt_env$i.txt)"
done
