#!/bin/bash
mkdir -p gpurun_out/c31
timeout 900 python -m pytest tests -m gpu -q -x -s -k "sweep_device" > gpurun_out/c31/tests.log 2>&1; echo tests=$?
grep -E "fp16 storage|caa.Net|passed|failed|FAILED|Error|^E " gpurun_out/c31/tests.log | head
