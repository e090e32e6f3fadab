#!/bin/bash
O=gpurun_out/c12
mkdir -p $O
timeout 600 python bench.py --steps 10 --warmup 3 > $O/bench_n1.json 2> $O/bench_n1.err; echo n1=$?
python - <<'PY'
import json
d=[json.loads(l) for l in open("gpurun_out/c12/bench_n1.json") if l.startswith("{")][-1]
print("value %.0f ms %.2f e2e %.0f frac %.3f"%(d["value"], d["ms_per_step"], d["e2e"]["value"], d["roofline"]["frac"]), d["clocks"])
PY
tail -3 $O/bench_n1.err
