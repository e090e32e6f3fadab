#!/bin/bash
O=gpurun_out/c28
mkdir -p $O
AY_L0_TS=1 timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "forward or layers or plan" > $O/tests.log 2>&1; echo tests_with_ts_default=$?
grep -E "passed|failed|FAILED|Error|^E " $O/tests.log | head -12
AY_L0_TS=1 timeout 300 python tools/layer_times.py --batch 256 --reps 5 --out $O/lt_ts.txt > $O/lt1.log 2>&1; echo lt1=$?
timeout 300 python tools/layer_times.py --batch 256 --reps 5 --out $O/lt_staged.txt > $O/lt2.log 2>&1; echo lt2=$?
echo "TS:"; head -3 $O/lt_ts.txt; grep -E "sum us|whole" $O/lt_ts.txt; echo "staged:"; head -3 $O/lt_staged.txt; grep -E "sum us|whole" $O/lt_staged.txt
