#!/bin/bash
mkdir -p gpurun_out/c19
timeout 600 python tools/next_rows_bench.py --out gpurun_out/c19/next_rows.txt 2>&1 | tail -12
