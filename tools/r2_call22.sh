#!/bin/bash
# 4-GPU check of both bench arms as the driver launches them
O=gpurun_out/c22
mkdir -p $O
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 4 --steps 10 --warmup 3"
timeout 600 $RUN --impl reference --steps 2 --warmup 1 > $O/ref_n4.json 2> $O/ref_n4.err; echo ref=$?
timeout 600 $RUN > $O/bench_n4.json 2> $O/bench_n4.err; echo n4=$?
python - <<'PY'
import json
for n in ("ref_n4", "bench_n4"):
    d=[json.loads(l) for l in open("gpurun_out/c22/%s.json" % n) if l.startswith("{")][-1]
    print(n, d.get("impl"), "value %.1f ms %.2f" % (d["value"], d["ms_per_step"]), "e2e", d.get("e2e", {}).get("value"), d.get("gather_verified"), d.get("per_rank", {}).get("ms_per_step"))
PY
