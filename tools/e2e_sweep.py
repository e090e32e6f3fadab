"""Sweep of detect()'s sub-batch plan for host input: python tools/e2e_sweep.py (GPU box)."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from amyloid_yolo_paper_b200 import synth
from amyloid_yolo_paper_b200.config import ensure_cfg
from amyloid_yolo_paper_b200.models import Darknet
from amyloid_yolo_paper_b200.pipeline import TileDetector

dev = torch.device("cuda", 0)
model = Darknet(ensure_cfg("yolov3-custom"))
synth.synthetic_state_dict(model, seed=0, conf_bias=(-6.0, -6.0, 0.62), wh_gain=0.15)
det = TileDetector(model, conf_thres=0.8, nms_thres=0.4, merge_boxes=True, max_batch=256, cap=128, device=dev)
B = 256
x = torch.rand((B, 3, 416, 416), dtype=torch.float32).pin_memory()
plans = [None, [32, 96, 96, 32], [16, 48, 64, 128], [32, 96, 112, 16], [16, 112, 112, 16], [64, 128, 64], [32, 64, 64, 64, 32], [128, 128]]
for plan in plans + plans[:1]:
    det.sub_batches = plan
    for _ in range(2):
        det.detect(x, (1536, 1536))
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(6):
        det.detect(x, (1536, 1536))
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / 6
    print("%-26s %7.2f ms  %8.0f tiles/s" % (plan, dt * 1e3, B / dt), flush=True)
