#!/bin/bash
O=gpurun_out/c7
mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -q -s > $O/tests.log 2>&1; echo tests=$?
grep -E "passed|failed|FAILED|fp32 mode, Gaussian|tf32x3" $O/tests.log | head -20
timeout 400 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > $O/bench_c3.json 2> $O/bench_c3.err; echo c3=$?
python - <<'PY'
import json
d=json.load(open("gpurun_out/c7/bench_c3.json"))
print("c3 value %.1f"%d["value"], "e2e %.1f"%d["e2e"]["value"], "fwd ms %.3f"%d["roofline"]["ms"], "frac %.4f"%d["roofline"]["frac"])
PY
