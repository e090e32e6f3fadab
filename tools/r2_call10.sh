#!/bin/bash
O=gpurun_out/c10
mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q -x -s -k "caa or evaluate_and" > $O/tests.log 2>&1; echo tests=$?
grep -E "caa.Net|CAA filter|prediction dictionary|passed|failed|FAILED|Error|assert|^E " $O/tests.log | head -40
