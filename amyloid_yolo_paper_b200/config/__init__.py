"""Network definitions, generated instead of shipped.

The reference keeps hand-edited darknet cfg files under ``config/``
(``config/yolov3-custom.cfg``, ``config/yolov3-tiny.cfg``; the custom one is
produced by ``config/create_custom_model.sh <num_classes>``).  This package
regenerates equivalent files from the architecture description so that the
GPU box (which has no copy of the reference) can build the same networks:

    path = ensure_cfg("yolov3-custom")          # 2 classes, 21-channel heads
    path = ensure_cfg("yolov3-tiny")            # 80 classes, 255-channel heads

``tests/test_cfg.py`` checks (in the build container, where the reference is
mounted) that the parsed block lists are identical to the reference's files.
"""
import os

_NET_TRAIN = [
    ("batch", "16"), ("subdivisions", "1"), ("width", "416"), ("height", "416"),
    ("channels", "3"), ("momentum", "0.9"), ("decay", "0.0005"), ("angle", "0"),
    ("saturation", "1.5"), ("exposure", "1.5"), ("hue", ".1"),
    ("learning_rate", "0.001"), ("burn_in", "1000"), ("max_batches", "500200"),
    ("policy", "steps"), ("steps", "400000,450000"), ("scales", ".1,.1"),
]

_ANCHORS_V3 = "10,13,  16,30,  33,23,  30,61,  62,45,  59,119,  116,90,  156,198,  373,326"
_ANCHORS_TINY = "10,14,  23,27,  37,58,  81,82,  135,169,  344,319"


def _block(kind, pairs):
    return "[%s]\n%s\n" % (kind, "\n".join("%s=%s" % kv for kv in pairs))


def _conv(filters, size, stride=1, bn=True, act="leaky"):
    pairs = []
    if bn:
        pairs.append(("batch_normalize", "1"))
    pairs += [("filters", str(filters)), ("size", str(size)), ("stride", str(stride)),
              ("pad", "1"), ("activation", act)]
    return _block("convolutional", pairs)


def _yolo(mask, anchors, classes, num):
    return _block("yolo", [("mask", mask), ("anchors", anchors), ("classes", str(classes)),
                           ("num", str(num)), ("jitter", ".3"), ("ignore_thresh", ".7"),
                           ("truth_thresh", "1"), ("random", "1")])


def yolov3_cfg_text(num_classes=2):
    """Darknet-53 + 3-scale YOLOv3 head (what create_custom_model.sh emits)."""
    head_filters = 3 * (5 + num_classes)
    out = [_block("net", _NET_TRAIN)]
    out.append(_conv(32, 3))
    for width, repeats in ((64, 1), (128, 2), (256, 8), (512, 8), (1024, 4)):
        out.append(_conv(width, 3, stride=2))
        for _ in range(repeats):
            out.append(_conv(width // 2, 1))
            out.append(_conv(width, 3))
            out.append(_block("shortcut", [("from", "-3"), ("activation", "linear")]))
    scales = ((512, "6,7,8", None), (256, "3,4,5", "61"), (128, "0,1,2", "36"))
    for narrow, mask, lateral in scales:
        if lateral is not None:
            out.append(_block("route", [("layers", "-4")]))
            out.append(_conv(narrow, 1))
            out.append(_block("upsample", [("stride", "2")]))
            out.append(_block("route", [("layers", "-1, " + lateral)]))
        for _ in range(3):
            out.append(_conv(narrow, 1))
            out.append(_conv(narrow * 2, 3))
        out.append(_conv(head_filters, 1, bn=False, act="linear"))
        out.append(_yolo(mask, _ANCHORS_V3, num_classes, 9))
    return "\n".join(out)


def yolov3_tiny_cfg_text(num_classes=80):
    head_filters = 3 * (5 + num_classes)
    net = [("batch", "1")] + _NET_TRAIN[1:]
    out = [_block("net", net)]
    for i, width in enumerate((16, 32, 64, 128, 256, 512)):
        out.append(_conv(width, 3))
        out.append(_block("maxpool", [("size", "2"), ("stride", "1" if i == 5 else "2")]))
    out.append(_conv(1024, 3))
    out.append(_conv(256, 1))
    out.append(_conv(512, 3))
    out.append(_conv(head_filters, 1, bn=False, act="linear"))
    out.append(_yolo("3,4,5", _ANCHORS_TINY, num_classes, 6))
    out.append(_block("route", [("layers", "-4")]))
    out.append(_conv(128, 1))
    out.append(_block("upsample", [("stride", "2")]))
    out.append(_block("route", [("layers", "-1, 8")]))
    out.append(_conv(256, 3))
    out.append(_conv(head_filters, 1, bn=False, act="linear"))
    out.append(_yolo("1,2,3", _ANCHORS_TINY, num_classes, 6))
    return "\n".join(out)


_GENERATORS = {
    "yolov3-custom": lambda: yolov3_cfg_text(2),
    "yolov3": lambda: yolov3_cfg_text(80),
    "yolov3-tiny": lambda: yolov3_tiny_cfg_text(80),
}


def ensure_cfg(name, directory=None):
    """Write ``<name>.cfg`` (if missing) and return its path."""
    if name not in _GENERATORS:
        raise KeyError("unknown network %r (have %s)" % (name, sorted(_GENERATORS)))
    directory = directory or os.path.join(os.path.dirname(os.path.abspath(__file__)), "_generated")
    os.makedirs(directory, exist_ok=True)
    path = os.path.join(directory, name + ".cfg")
    text = _GENERATORS[name]()
    if not os.path.exists(path) or open(path).read() != text:
        tmp = path + ".tmp%d" % os.getpid()
        with open(tmp, "w") as fh:
            fh.write(text)
        os.replace(tmp, path)
    return path
