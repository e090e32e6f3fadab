#!/bin/bash
mkdir -p gpurun_out/c31
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -s -k "fp16" > gpurun_out/c31/tests.log 2>&1; echo tests=$?
grep -E "fp16 storage|passed|failed|FAILED|Error|^E " gpurun_out/c31/tests.log | head
