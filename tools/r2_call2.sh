#!/bin/bash
O=gpurun_out/c2
mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -q -s -x -k "candidate or pipeline_end or fp32_mode_gaussian or detect_stream or jpeg or planted" > $O/tests.log 2>&1; echo tests=$?
tail -8 $O/tests.log
timeout 400 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > $O/bench_c3.json 2> $O/bench_c3.err; echo c3=$?
timeout 300 python bench.py --config c1 --steps 30 --warmup 5 --no-cpu-baseline > $O/bench_c1.json 2> $O/bench_c1.err; echo c1=$?
python - <<'PY'
import json
for c in ("c3","c1"):
    d=json.load(open("gpurun_out/c2/bench_%s.json"%c))
    print(c, "value %.1f"%d["value"], "ms/step %.3f"%d["ms_per_step"], "e2e %.1f"%d["e2e"]["value"], "fwd ms %.3f"%d["roofline"]["ms"], "launches", d["gpu_launches"])
PY
