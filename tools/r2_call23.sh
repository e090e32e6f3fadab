#!/bin/bash
O=gpurun_out/c23
mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_bars.py -m gpu -q -x > $O/tests.log 2>&1; echo tests=$?
grep -E "passed|failed|FAILED|Error|^E " $O/tests.log | head
timeout 300 python tools/layer_times.py --batch 256 --reps 5 --counters --out $O/lt.txt > $O/lt.log 2>&1; echo lt=$?
grep -E "^ *(0|1|2|3|5|7|10|14|17|35|39|64|100) " $O/lt.txt | head -30; grep -E "sum us|whole" $O/lt.txt
timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > $O/bench.json 2> $O/b.err; echo b=$?
python - <<'PY'
import json
d=[json.loads(l) for l in open("gpurun_out/c23/bench.json") if l.startswith("{")][-1]
print("value %.0f ms %.2f e2e %.0f conv ms %.3f frac %.4f" % (d["value"], d["ms_per_step"], d["e2e"]["value"], d["roofline"]["ms"], d["roofline"]["frac"]))
PY
