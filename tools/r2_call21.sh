#!/bin/bash
mkdir -p gpurun_out/c21
timeout 900 python -m pytest tests -m gpu -q -x -k "tiny or layers or mini or B64 or batch or candidate or plan_variants or pipeline" > gpurun_out/c21/tests.log 2>&1; echo tests=$?
grep -E "passed|failed|FAILED|Error|^E " gpurun_out/c21/tests.log | head
timeout 300 python tools/layer_times.py --net yolov3-tiny --batch 64 --reps 5 --out gpurun_out/c21/lt_tiny.txt > gpurun_out/c21/log.txt 2>&1; echo rc=$?
cat gpurun_out/c21/lt_tiny.txt | head -28; grep -E "sum us|whole" gpurun_out/c21/lt_tiny.txt
timeout 300 python bench.py --config c2 --steps 20 --warmup 3 > gpurun_out/c21/bench_c2.json 2> gpurun_out/c21/c2.err; echo c2=$?
python - <<'PY'
import json
d=[json.loads(l) for l in open("gpurun_out/c21/bench_c2.json") if l.startswith("{")][-1]
print("c2 value %.0f ms %.3f e2e %.0f frac %.4f" % (d["value"], d["ms_per_step"], d["e2e"]["value"], d["roofline"]["frac"]))
PY
