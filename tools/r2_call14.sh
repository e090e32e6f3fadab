#!/bin/bash
O=gpurun_out/c14
mkdir -p $O
timeout 600 python -m pytest tests -m gpu -q -x -k "caa or test_gpu_parity" > $O/tests.log 2>&1; echo tests=$?
grep -E "passed|failed|FAILED|Error|^E " $O/tests.log | head
timeout 300 python tools/layer_times.py --batch 256 --reps 5 --counters --out $O/lt.txt > $O/lt.log 2>&1; echo lt=$?
head -12 $O/lt.txt; grep -A1 "cycle counters" $O/lt.txt | tail -1; grep -E "sum us|whole" $O/lt.txt
timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > $O/bench.json 2> $O/b.err; echo b=$?
python - <<'PY'
import json
d=[json.loads(l) for l in open("gpurun_out/c14/bench.json") if l.startswith("{")][-1]
print("value %.0f ms %.2f e2e %.0f conv ms %.3f frac %.4f" % (d["value"], d["ms_per_step"], d["e2e"]["value"], d["roofline"]["ms"], d["roofline"]["frac"]), d["clocks"])
PY
