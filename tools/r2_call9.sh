#!/bin/bash
# the allocation-free NMS entry point: alone in a fresh process, then after the whole GPU suite
O=gpurun_out/c9
mkdir -p $O
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -x > $O/alone.log 2>&1; echo alone=$?
tail -15 $O/alone.log
timeout 2400 python -m pytest tests -m gpu -q -x -s > $O/tests.log 2>&1; echo tests=$?
grep -E "passed|failed|FAILED|Error|^E " $O/tests.log | head -30
