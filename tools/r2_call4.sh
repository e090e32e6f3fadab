#!/bin/bash
O=gpurun_out/c4
mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q -x -k "batch_statistics or evaluate_and" > $O/tests.log 2>&1; echo tests=$?
grep -E "passed|failed|FAILED|Error|assert|^E " $O/tests.log | head -30
