#!/bin/bash
O=gpurun_out/c10
mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q -x -s -k "caa" > $O/tests.log 2>&1; echo tests=$?
grep -E "caa.Net|CAA filter|slide|passed|failed|FAILED|Error|assert|^E " $O/tests.log | head -40
timeout 600 python tools/next_rows_bench.py 2>&1 | grep "N2" | head -4
