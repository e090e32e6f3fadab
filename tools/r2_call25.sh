#!/bin/bash
O=gpurun_out/c25
mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "forward or layers or mini or plan" > $O/tests.log 2>&1; echo tests=$?
grep -E "passed|failed|FAILED|Error|^E " $O/tests.log | head
timeout 300 python tools/layer_times.py --batch 256 --reps 5 --out $O/lt.txt > $O/lt.log 2>&1; echo lt=$?
head -4 $O/lt.txt; grep -E "sum us|whole" $O/lt.txt
