#!/bin/bash
O=gpurun_out/c29
mkdir -p $O
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -x > $O/alone.log 2>&1; echo alone=$?
timeout 2400 python -m pytest tests -m gpu -q -x -s > $O/tests.log 2>&1; echo tests=$?
grep -E "passed|failed|FAILED|Error|^E " $O/tests.log | head -20
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo smoke=$?; tail -1 $O/smoke.log
timeout 400 python bench.py --steps 20 --warmup 3 > $O/bench_c3.json 2> $O/b.err; echo c3=$?
python - <<'PY'
import json
d=[json.loads(l) for l in open("gpurun_out/c29/bench_c3.json") if l.startswith("{")][-1]
print("value %.0f ms %.2f e2e %.0f frac %.4f" % (d["value"], d["ms_per_step"], d["e2e"]["value"], d["roofline"]["frac"]), d["cpu_baseline"]["value"], d["gpu_launches"])
PY
