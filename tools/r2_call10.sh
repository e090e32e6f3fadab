#!/bin/bash
O=gpurun_out/c10
mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q -x -s -k "caa or slide" > $O/tests.log 2>&1; echo tests=$?
grep -E "caa.Net|CAA filter|slide|passed|failed|FAILED|Error|assert|^E " $O/tests.log | head -40
