#!/bin/bash
O=gpurun_out/c4
mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q -s -x -k "fp32_mode_gaussian" > $O/tests.log 2>&1; echo tests=$?
grep -E "passed|failed|FAILED|Error|fp32 mode|assert" $O/tests.log | head -20
