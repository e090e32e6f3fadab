#!/bin/bash
# Ablation sweep of the conv GEMM (AY_DBG bits: 1 no MMA, 2 no operand TMA, 4 no output store, 8 no residual).
# usage (on the GPU box): bash tools/ablate.sh "0 1 2 4" 64
for D in ${1:-0 1 2 4 8}; do
AY_DBG=$D timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"conv_gemm|first_conv|prep_input|s2_gather|upsample2" -c 82 --csv --log-file gpurun_out/abl_$D.csv python bench.py --steps 1 --warmup 1 --batch ${2:-64} --no-cpu-baseline > gpurun_out/abl_$D.log 2>&1
done
