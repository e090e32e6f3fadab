#!/bin/bash
# first layer with paired staged pixels (one K=16 MMA per filter row): parity, then A/B of the per-layer times
O=gpurun_out/c13
mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x > $O/tests.log 2>&1; echo tests=$?
grep -E "passed|failed|FAILED|Error|^E " $O/tests.log | head -20
timeout 300 python tools/layer_times.py --batch 256 --reps 5 --counters --out $O/lt_paired.txt > $O/lt1.log 2>&1; echo lt1=$?
AY_L0_UNPAIRED=1 timeout 300 python tools/layer_times.py --batch 256 --reps 5 --counters --out $O/lt_unpaired.txt > $O/lt2.log 2>&1; echo lt2=$?
echo "layer | paired | unpaired"
paste <(head -12 $O/lt_paired.txt | cut -c1-44) <(head -12 $O/lt_unpaired.txt | cut -c33-44)
grep -A2 "cycle counters" $O/lt_paired.txt | tail -1; grep -A2 "cycle counters" $O/lt_unpaired.txt | tail -1
grep -iE "sum|whole|prep" $O/lt_paired.txt $O/lt_unpaired.txt | head
timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > $O/bench_paired.json 2> $O/b1.err; echo b1=$?
AY_L0_UNPAIRED=1 timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > $O/bench_unpaired.json 2> $O/b2.err; echo b2=$?
python - <<'PY'
import json
for n in ("paired", "unpaired"):
    d=[json.loads(l) for l in open("gpurun_out/c13/bench_%s.json" % n) if l.startswith("{")][-1]
    print(n, "value %.0f ms %.2f e2e %.0f conv ms %.3f frac %.4f" % (d["value"], d["ms_per_step"], d["e2e"]["value"], d["roofline"]["ms"], d["roofline"]["frac"]), d["clocks"])
PY
