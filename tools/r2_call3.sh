#!/bin/bash
O=gpurun_out/c3
mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -q -s > $O/tests.log 2>&1; echo tests=$?
grep -E "passed|failed|FAILED|Error" $O/tests.log | head -20
timeout 400 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > $O/bench_c3.json 2> $O/bench_c3.err; echo c3=$?
timeout 200 python tools/attribution.py --out $O/attribution.txt > $O/attr.log 2>&1; echo attr=$?
timeout 300 python tools/layer_times.py --batch 256 --reps 4 --out $O/layer_times.txt > $O/lt.log 2>&1; echo lt=$?
python - <<'PY'
import json
d=json.load(open("gpurun_out/c3/bench_c3.json"))
print("c3 value %.1f"%d["value"], "ms/step %.3f"%d["ms_per_step"], "e2e %.1f"%d["e2e"]["value"], "fwd ms %.3f"%d["roofline"]["ms"], "frac %.4f"%d["roofline"]["frac"])
PY
tail -10 $O/attribution.txt; grep -E "sum us|whole" $O/layer_times.txt
