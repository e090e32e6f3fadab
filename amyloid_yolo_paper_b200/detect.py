"""``detect.py`` with the reference's command line (``detect.py:30-47``), headless.

    python -m amyloid_yolo_paper_b200.detect --image_folder tiles/ --model_def <cfg> \
        --weights_path checkpoints/yolov3_ckpt_105.pth --class_path data/custom/classes.names \
        --conf_thres 0.8 --nms_thres 0.4 --img_size 416 --batch_size 64 --n_cpu 8 --merge_boxes True

Same flags, defaults and string-boolean convention (``--merge_boxes`` is compared with ``"True"``,
``detect.py:132``).  ``--img_size`` is the network input size exactly as in the reference (``Darknet(...,
img_size)``, ``Resize(opt.img_size)``, ``rescale_boxes(..., opt.img_size, ...)``: detect.py:59,73,129), any multiple
of 32; ``--batch_size`` and ``--n_cpu`` size the batches and the decode workers of the loader
(``utils/datasets.py``), which hands pinned uint8 blocks of the 416^2 sampled pixels to one
H2D / compute / D2H pipeline (``TileDetector.detect_stream``).

Differences, all outside the numeric path: nothing is drawn (the reference's matplotlib output,
``detect.py:108-171``, is left out); detections are printed like the reference prints them and written to
``<output>/detections.json``; the CAA-classifier flags are accepted and rejected with a clear message (out of
scope).
"""
import argparse
import json
import os
import time

import numpy as np
import torch

from .models import Darknet
from .pipeline import TileDetector
from .utils.datasets import TileLoader, load_image  # noqa: F401  (load_image re-exported for callers of round 1)
from .utils.utils import load_classes


def build_parser():
    p = argparse.ArgumentParser()
    p.add_argument("--image_folder", type=str, default="data/samples", help="path to dataset")
    p.add_argument("--model_def", type=str, default="config/yolov3.cfg", help="path to model definition file")
    p.add_argument("--weights_path", type=str, default="weights/yolov3.weights", help="path to weights file")
    p.add_argument("--class_path", type=str, default="data/coco.names", help="path to class label file")
    p.add_argument("--conf_thres", type=float, default=0.8, help="object confidence threshold")
    p.add_argument("--nms_thres", type=float, default=0.4, help="iou thresshold for non-maximum suppression")
    p.add_argument("--batch_size", type=int, default=1, help="size of the batches")
    p.add_argument("--n_cpu", type=int, default=0, help="number of cpu threads to use during batch generation")
    p.add_argument("--img_size", type=int, default=416, help="size of each image dimension")
    p.add_argument("--checkpoint_model", type=str, help="path to checkpoint model")
    p.add_argument("--merge_boxes", type=str, help="whether to merge bboxes of the same classes that are overlapping")
    p.add_argument("--write_CAA_detections_to_pickle", type=str, help="(reference flag; unsupported here)")
    p.add_argument("--filter_CAA_detections_by_model", type=str, help="(reference flag; unsupported here)")
    p.add_argument("--output", type=str, default="output", help="directory for detections.json")
    return p


def pad_square(img):
    """utils/transforms.py:80-86 (imgaug PadToAspectRatio(1.0, position='center-center'),
    constant 0): centred zero padding to a square; identity for square tiles."""
    h, w = img.shape[:2]
    if h == w:
        return img
    s = max(h, w)
    out = np.zeros((s, s, 3), dtype=np.uint8)
    top, left = (s - h) // 2, (s - w) // 2
    out[top:top + h, left:left + w] = img
    return out


def load_model(opt, device):
    model = Darknet(opt.model_def, img_size=opt.img_size).to(device)
    if opt.weights_path.endswith(".weights"):
        model.load_darknet_weights(opt.weights_path)
    else:
        model.load_state_dict(torch.load(opt.weights_path, map_location="cpu"))
    model.eval()
    return model


def run(opt):
    if opt.write_CAA_detections_to_pickle == "True" or opt.filter_CAA_detections_by_model == "True":
        raise SystemExit("the CAA-classifier post-filter (core.py:425-480) is outside this build's scope")
    if not torch.cuda.is_available():
        raise SystemExit("no CUDA device: this detect.py runs on the B200 engine only (no CPU fallback)")
    if opt.img_size < 32 or opt.img_size % 32 != 0:
        raise SystemExit("--img_size must be a positive multiple of 32 (got %d)" % opt.img_size)
    device = torch.device("cuda", torch.cuda.current_device())
    os.makedirs(opt.output, exist_ok=True)
    model = load_model(opt, device)
    classes = load_classes(opt.class_path) if opt.class_path and os.path.exists(opt.class_path) else None
    batch = max(1, opt.batch_size)
    det = TileDetector(model, conf_thres=opt.conf_thres, nms_thres=opt.nms_thres, merge_boxes=(opt.merge_boxes == "True"),
                       max_batch=batch, device=device, img_size=opt.img_size)
    loader = TileLoader(opt.image_folder, img_size=opt.img_size, batch_size=batch, n_cpu=opt.n_cpu)
    results, order = {}, []

    def batches():                                # (pinned uint8 [b,S,S,3], original (h, w)) per loader batch
        for paths, tiles, shape in loader:
            order.append(paths)
            yield tiles, shape

    print("\nPerforming object detection:")
    prev = time.time()
    for bi, outs in enumerate(det.detect_stream(batches())):
        for p, out in zip(order[bi], outs):
            results[p] = [] if out is None else out.tolist()
        now = time.time()
        print("\t+ Batch %d, Inference Time: %.4fs" % (bi, now - prev))
        prev = now
    for p, rows in results.items():
        print("Image: '%s'" % p)
        for x1, y1, x2, y2, conf, cls_conf, cls_pred in rows:
            name = classes[int(cls_pred)] if classes and int(cls_pred) < len(classes) else str(int(cls_pred))
            print("\t+ Label: %s, Conf: %.5f" % (name, cls_conf))
    with open(os.path.join(opt.output, "detections.json"), "w") as fh:
        json.dump(results, fh)
    return results


def main(argv=None):
    opt = build_parser().parse_args(argv)
    print(opt)
    run(opt)


if __name__ == "__main__":
    main()
