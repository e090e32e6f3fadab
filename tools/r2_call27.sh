#!/bin/bash
O=gpurun_out/c27
mkdir -p $O
timeout 600 python bench.py --steps 100 --warmup 5 --no-cpu-baseline > $O/bench_k100.json 2> $O/b.err; echo b=$?
python - <<'PY'
import json
d=[json.loads(l) for l in open("gpurun_out/c27/bench_k100.json") if l.startswith("{")][-1]
print("K=100: value %.0f ms %.2f e2e %.0f conv ms %.3f frac %.4f" % (d["value"], d["ms_per_step"], d["e2e"]["value"], d["roofline"]["ms"], d["roofline"]["frac"]), d["clocks"])
PY
