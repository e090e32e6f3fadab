#!/bin/bash
O=gpurun_out/c8
mkdir -p $O
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo smoke=$?; tail -2 $O/smoke.log
timeout 1500 python -m pytest tests -m gpu -q -x > $O/tests.log 2>&1; echo tests=$?
grep -E "passed|failed|FAILED|Error" $O/tests.log | head
timeout 300 python tools/layer_times.py --batch 256 --reps 4 --out $O/lt_tma.txt > $O/lt1.log 2>&1
AY_NO_TMASTORE=1 timeout 300 python tools/layer_times.py --batch 256 --reps 4 --out $O/lt_notma.txt > $O/lt2.log 2>&1
paste <(awk '{print $1,$2,$3,$4,$5,$6,$7}' $O/lt_tma.txt | head -45) <(awk '{print $7}' $O/lt_notma.txt | head -45)
grep -E "sum us|whole" $O/lt_tma.txt $O/lt_notma.txt
