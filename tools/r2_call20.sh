#!/bin/bash
mkdir -p gpurun_out/c20
timeout 300 python tools/attribution.py --batch 1 --heat 0.5 --out gpurun_out/c20/attr_b1.txt > gpurun_out/c20/log.txt 2>&1; echo rc=$?
head -14 gpurun_out/c20/attr_b1.txt; tail -14 gpurun_out/c20/attr_b1.txt
AY_NO_PDL=1 timeout 300 python tools/attribution.py --batch 1 --heat 0.5 --out gpurun_out/c20/attr_b1_nopdl.txt > gpurun_out/c20/log2.txt 2>&1
tail -8 gpurun_out/c20/attr_b1_nopdl.txt
