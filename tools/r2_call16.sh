#!/bin/bash
# final pass of the round: whole GPU suite (the allocation-free NMS test alone first), smoke, then the measurement pass
O=gpurun_out/c16
mkdir -p $O
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -x > $O/alone.log 2>&1; echo alone=$?
timeout 2400 python -m pytest tests -m gpu -q -x -s > $O/tests.log 2>&1; echo tests=$?
grep -E "passed|failed|FAILED|Error|^E " $O/tests.log | head -20
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo smoke=$?; tail -1 $O/smoke.log
bash tools/r2_measure.sh > $O/measure.log 2>&1; echo measure=$?
grep -E "^[a-z0-9_]+=[0-9]+$" $O/measure.log | tr '\n' ' '
