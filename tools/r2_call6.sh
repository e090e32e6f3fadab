#!/bin/bash
O=gpurun_out/c6
mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q -x -k "fp32" > $O/tests.log 2>&1; echo tests=$?; tail -2 $O/tests.log
timeout 300 python bench.py --precision fp32 --batch 64 --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > $O/bench_f32x3.json 2> $O/f32x3.err; echo f32x3=$?
AY_F32_FFMA=1 timeout 600 python bench.py --precision fp32 --batch 64 --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > $O/bench_f32ffma.json 2> $O/f32ffma.err; echo ffma=$?
python - <<'PY'
import json
for n in ("f32x3","f32ffma"):
    try:
        d=json.load(open("gpurun_out/c6/bench_%s.json"%n))
        print(n, d["dtype"], "value %.1f tiles/s"%d["value"], "fwd ms %.2f"%d["roofline"]["ms"], "TF %.1f"%d["roofline"]["achieved"], "frac %.3f"%d["roofline"]["frac"])
    except Exception as e: print(n, "failed", e)
PY
tail -3 $O/f32x3.err $O/f32ffma.err
