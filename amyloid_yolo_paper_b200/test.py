"""``evaluate`` with the reference's signature (``test.py:24-66``) and the prediction dictionary of
``prospective.runModelOnValidationImages`` (``prospective.py:32-93``) on top of the B200 engine.

``evaluate`` walks a list file of image paths (labels beside them: ``images`` -> ``labels``, ``.jpg/.png`` -> ``.txt``,
rows ``class x y w h`` normalised, ``utils/datasets.py:64-105``), runs forward + NMS on the GPU in network-input pixels
(no rescale, no merge - as ``test.py:52-57``), matches predictions to labels (``get_batch_statistics``) and returns
``(precision, recall, AP, f1, ap_class)`` from ``ap_per_class``.  Square images only (the slide tiles): the
reference's label path through imgaug's ``PadSquare`` is the identity there, up to clipping boxes to the image, which is
done here as well.
"""
import numpy as np
import torch

from .pipeline import TileDetector
from .utils.datasets import TileLoader, gather_resized, load_image
from .utils.utils import ap_per_class, get_batch_statistics, non_max_suppression_device, xywh2xyxy


def _load_labels(image_path):
    label_path = image_path.replace("images", "labels").replace(".png", ".txt").replace(".jpg", ".txt")
    try:
        import warnings
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            boxes = np.loadtxt(label_path).reshape(-1, 5)
    except Exception:
        print("Could not read label '%s'." % label_path)
        return None
    # imgaug's clip_out_of_image on the absolute xyxy box, back to normalised xywh (utils/transforms.py:19-54)
    x1, y1 = np.clip(boxes[:, 1] - boxes[:, 3] / 2, 0, 1), np.clip(boxes[:, 2] - boxes[:, 4] / 2, 0, 1)
    x2, y2 = np.clip(boxes[:, 1] + boxes[:, 3] / 2, 0, 1), np.clip(boxes[:, 2] + boxes[:, 4] / 2, 0, 1)
    out = np.stack([boxes[:, 0], (x1 + x2) / 2, (y1 + y2) / 2, x2 - x1, y2 - y1], 1)
    return out[(out[:, 3] > 0) & (out[:, 4] > 0)]


def evaluate(model, path, iou_thres, conf_thres, nms_thres, img_size, batch_size):
    """-> (precision, recall, AP, f1, ap_class), or None when nothing was detected over the whole set."""
    model.eval()
    with open(path, "r") as fh:
        files = [line.rstrip() for line in fh if line.strip()]
    device = torch.device("cuda", torch.cuda.current_device())
    eng = model.engine(device=device, max_batch=batch_size, img_size=img_size)
    labels, sample_metrics = [], []
    for b0 in range(0, len(files), batch_size):
        imgs, targets = [], []
        for f in files[b0:b0 + batch_size]:
            boxes = _load_labels(f)
            if boxes is None:
                continue
            img = load_image(f)
            if img.shape[0] != img.shape[1]:
                raise ValueError("evaluate() handles square tiles only (got %s for %s)" % (img.shape[:2], f))
            t = torch.zeros((len(boxes), 6))
            t[:, 1:] = torch.from_numpy(boxes).to(torch.float32)
            t[:, 0] = len(imgs)
            imgs.append(torch.from_numpy(gather_resized(img, img_size)))
            targets.append(t)
        if not imgs:
            continue
        targets = torch.cat(targets, 0)
        labels += targets[:, 1].tolist()
        targets[:, 2:] = xywh2xyxy(targets[:, 2:])
        targets[:, 2:] *= img_size
        decoded = eng.forward_u8(torch.stack(imgs))
        rows, _, counts = non_max_suppression_device(decoded, conf_thres, nms_thres)
        eng.check()
        counts = counts.cpu().tolist()
        rows = rows[:, :max(max(counts), 1)].cpu()
        outputs = [rows[i, :k].clone() if k > 0 else None for i, k in enumerate(counts)]
        sample_metrics += get_batch_statistics(outputs, targets, iou_threshold=iou_thres)
    if len(sample_metrics) == 0:
        return None
    true_positives, pred_scores, pred_labels = [np.concatenate([np.asarray(v) for v in x], 0) for x in zip(*sample_metrics)]
    return ap_per_class(true_positives, pred_scores, pred_labels, labels)


def run_model_on_validation_images(model, image_folder, classes, conf_thres=0.8, nms_thres=0.4, img_size=416, batch_size=8,
                                   n_cpu=12, merge_boxes=True, strip_prefix=None, caa_model=None, caa_norm=None):
    """The dictionary ``prospective.runModelOnValidationImages`` pickles (``prospective.py:84-92``): image name ->
    list of ``({'x1','x2','y1','y2','conf','cls_conf','cls_pred'}, class name)``; images without detections map to an
    empty list.  With ``caa_model`` (a ``caa.Net``) and ``caa_norm`` the consensus CAA filter runs on every image's merged
    detections as in ``prospective.py:86-87`` (rows then come in the filter's set order, as in the reference)."""
    device = torch.device("cuda", torch.cuda.current_device())
    det = TileDetector(model, conf_thres=conf_thres, nms_thres=nms_thres, merge_boxes=merge_boxes, max_batch=batch_size,
                       device=device, img_size=img_size)
    loader = TileLoader(image_folder, img_size=img_size, batch_size=batch_size, n_cpu=n_cpu)
    order, out = [], {}

    def batches():
        for paths, tiles, shape in loader:
            order.append(paths)
            yield tiles, shape

    for bi, dets in enumerate(det.detect_stream(batches())):
        for p, d in zip(order[bi], dets):
            key = p.replace(strip_prefix if strip_prefix is not None else image_folder, "")
            out[key] = []
            if d is None:
                continue
            if caa_model is not None:
                from .caa import filterDetectionsByCAAModel
                d = filterDetectionsByCAAModel(p, d, classes, model=caa_model, norm=caa_norm, device=device)
                if len(d) == 0:                           # (everything filtered: prospective.py:88-89)
                    continue
            for x1, y1, x2, y2, conf, cls_conf, cls_pred in d:
                out[key].append(({"x1": x1.item(), "x2": x2.item(), "y1": y1.item(), "y2": y2.item(), "conf": conf.item(),
                                  "cls_conf": cls_conf.item(), "cls_pred": cls_pred.item()}, classes[int(cls_pred)]))
    return out
