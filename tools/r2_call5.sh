#!/bin/bash
O=gpurun_out/c5
mkdir -p $O
for cfg in "1 128" "2 128" "4 64" "2 64"; do
  set -- $cfg
  for k in 0.5 0.75 1.0; do
    AY_TF32_NA=$1 AY_TF32_BN=$2 AY_TF32_KAPPA=$k timeout 600 python -m pytest tests -m gpu -q -s -x -k "fp32_mode_full_networks and custom" > $O/t_$1_$2_$k.log 2>&1
    echo "NA=$1 BN=$2 kappa=$k rc=$? $(grep -E 'fp32 mode, head at layer 81' $O/t_$1_$2_$k.log | cut -c30-200)"
  done
done
