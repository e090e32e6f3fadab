#!/bin/bash
O=gpurun_out/c17
mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q -x -k "nms or iou or pipeline or no_alloc" > $O/tests.log 2>&1; echo tests=$?
grep -E "passed|failed|FAILED|Error|^E " $O/tests.log | head -20
timeout 400 python bench.py --config c5 --steps 5 --warmup 3 > $O/bench_c5.json 2> $O/c5.err; echo c5=$?
python - <<'PY'
import json
d=[json.loads(l) for l in open("gpurun_out/c17/bench_c5.json") if l.startswith("{")][-1]
print("c5 value %.0f ms %.2f e2e %.0f" % (d["value"], d["ms_per_step"], d["e2e"]["value"]), d["roofline"]["kernels_ms"], "frac %.4f" % d["roofline"]["frac"])
PY
