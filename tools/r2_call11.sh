#!/bin/bash
# 8-GPU scaling check of the default bench (config C4): device-timed value, e2e, gather_verified; with / without the
# collective warm-up, phases of the sweep on stderr
O=gpurun_out/c11
mkdir -p $O
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --steps 10 --warmup 3"
AY_SWEEP_TIMING=1 AY_NO_WARM_GATHER=1 timeout 600 $RUN --no-e2e --no-cpu-baseline > $O/bench_n8_nowarm.json 2> $O/bench_n8_nowarm.err; echo nowarm=$?
grep "\[sweep\]" $O/bench_n8_nowarm.err | tail -12
AY_SWEEP_TIMING=1 timeout 600 $RUN --no-e2e --no-cpu-baseline > $O/bench_n8_warm.json 2> $O/bench_n8_warm.err; echo warm=$?
grep "\[sweep\]" $O/bench_n8_warm.err | tail -12
timeout 600 $RUN > $O/bench_n8.json 2> $O/bench_n8.err; echo n8=$?
python - <<'PY'
import json
for n in ("bench_n8_nowarm", "bench_n8_warm", "bench_n8"):
    d=[json.loads(l) for l in open("gpurun_out/c11/%s.json" % n) if l.startswith("{")][-1]
    print(n, "value %.0f ms %.2f frac %.3f" % (d["value"], d["ms_per_step"], d["roofline"]["frac"]), d.get("e2e", {}).get("value"), d["per_rank"]["ms_per_step"][:3])
PY
