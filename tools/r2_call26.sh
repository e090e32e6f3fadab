#!/bin/bash
O=gpurun_out/c26
mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q -x -k "fp32 or f32 or caa" > $O/tests.log 2>&1; echo tests=$?
grep -E "passed|failed|FAILED|Error|^E " $O/tests.log | head
timeout 400 python tools/layer_times.py --batch 64 --reps 3 --precision fp32 --out $O/lt_f32.txt > $O/lt.log 2>&1; echo lt=$?
grep -E "^ *(0|1|2|3|4|5|6|12|13|37|38|62|63) " $O/lt_f32.txt; grep -E "sum us|whole" $O/lt_f32.txt
