"""Host-side logic that needs no GPU: cfg parsing, module/parameter naming, weight-file
formats, the C-ABI library's exports, loud failure without a device, CLI flags."""
import ctypes
import os
import re
import tempfile

import numpy as np
import pytest
import torch

from helpers import ROOT, write_mini_cfg
from amyloid_yolo_paper_b200 import _lib, synth
from amyloid_yolo_paper_b200.config import ensure_cfg
from amyloid_yolo_paper_b200.models import Darknet
from amyloid_yolo_paper_b200.utils.parse_config import parse_data_config, parse_model_config


def test_generated_cfgs_equal_reference_blocks(golden):
    """config/__init__.py regenerates the reference's cfg files; the parsed block lists must be
    identical to the reference parser's output on the reference's files (tests/golden/cfg_blocks.json)."""
    want = golden("cfg_blocks.json")
    for name in ("yolov3-custom", "yolov3-tiny", "yolov3"):
        got = parse_model_config(ensure_cfg(name))
        ref = want[name]
        assert len(got) == len(ref), name
        for i, (a, b) in enumerate(zip(got[1:], ref[1:])):
            a = {k: v for k, v in a.items()}
            b = {k: v for k, v in b.items()}
            if a["type"] == "yolo":                       # whitespace inside the anchor list differs harmlessly
                for k in ("anchors", "mask"):
                    a[k] = a[k].replace(" ", "")
                    b[k] = b[k].replace(" ", "")
            assert a == b, (name, i, a, b)
        for k in ("channels", "height", "width"):
            assert got[0][k] == ref[0][k]


def test_parser_quirks(tmp_path):
    p = tmp_path / "q.cfg"
    p.write_text("[net]\nheight=64\n# comment\n\n[convolutional]\nfilters = 8 \nsize=1\n[convolutional]\nbatch_normalize=1\nfilters=4\n")
    blocks = parse_model_config(str(p))
    assert blocks[1]["batch_normalize"] == 0 and blocks[1]["filters"] == "8"      # int 0 default, strings otherwise
    assert blocks[2]["batch_normalize"] == "1"
    d = tmp_path / "x.data"
    d.write_text("classes= 3\ntrain=a.txt\n")
    opt = parse_data_config(str(d))
    assert opt["classes"] == "3" and opt["gpus"] == "0,1,2,3" and opt["num_workers"] == "10"


def test_custom_network_layout_and_state_dict_keys():
    m = Darknet(ensure_cfg("yolov3-custom"))
    assert len(m.module_defs) == 107 and len(m.module_list) == 107 and len(m.yolo_layers) == 3
    sd = m.state_dict()
    assert len(sd) == 438
    assert "module_list.0.conv_0.weight" in sd and "module_list.0.batch_norm_0.running_var" in sd
    assert "module_list.81.conv_81.bias" in sd and "module_list.105.conv_105.bias" in sd
    assert sd["module_list.105.conv_105.weight"].shape == (21, 256, 1, 1)
    assert [y.anchors for y in m.yolo_layers] == [[(116, 90), (156, 198), (373, 326)], [(30, 61), (62, 45), (59, 119)],
                                                  [(10, 13), (16, 30), (33, 23)]]
    assert m.img_size == 416 and m.seen == 0 and m.header_info.tolist() == [0, 0, 0, 0, 0]


def test_param_count_matches_survey():
    m = Darknet(ensure_cfg("yolov3-custom"))
    total = sum(v.numel() for k, v in m.state_dict().items() if "num_batches_tracked" not in k)
    assert total == 61581727
    t = Darknet(ensure_cfg("yolov3-tiny"))
    total = sum(v.numel() for k, v in t.state_dict().items() if "num_batches_tracked" not in k)
    assert total == 8858734


def test_pth_roundtrip_and_darknet_weights_roundtrip(tmp_path):
    cfg = write_mini_cfg(tmp_path)
    a = Darknet(cfg, img_size=64)
    synth.synthetic_state_dict(a, seed=3)
    pth = str(tmp_path / "ckpt.pth")
    torch.save(a.state_dict(), pth)
    b = Darknet(cfg, img_size=64)
    missing = b.load_state_dict(torch.load(pth))
    assert not missing.missing_keys and not missing.unexpected_keys
    a.seen = 12345
    wpath = str(tmp_path / "mini.weights")
    a.save_darknet_weights(wpath)
    n_float = sum(v.numel() for k, v in a.state_dict().items() if "num_batches_tracked" not in k)
    assert os.path.getsize(wpath) == 20 + 4 * n_float                             # int32[5] header + fp32 stream
    c = Darknet(cfg, img_size=64)
    c.load_darknet_weights(wpath)
    assert int(c.seen) == 12345
    for (k, v), (_, w) in zip(a.state_dict().items(), c.state_dict().items()):
        if "num_batches_tracked" not in k:
            assert torch.equal(v, w), k
    # stream order per conv: bn.bias, bn.weight, running_mean, running_var, conv.weight (models.py:276-300)
    blob = np.fromfile(wpath, dtype=np.float32, offset=20)
    sd = a.state_dict()
    assert np.array_equal(blob[:32], sd["module_list.0.batch_norm_0.bias"].numpy())
    assert np.array_equal(blob[32:64], sd["module_list.0.batch_norm_0.weight"].numpy())
    assert np.array_equal(blob[128:128 + 864], sd["module_list.0.conv_0.weight"].numpy().reshape(-1))


def test_tiny_weights_file_size():
    """save_darknet_weights on yolov3-tiny is 35 434 956 B = the real yolov3-tiny.weights size."""
    t = Darknet(ensure_cfg("yolov3-tiny"))
    with tempfile.TemporaryDirectory() as d:
        p = os.path.join(d, "t.weights")
        t.save_darknet_weights(p)
        assert os.path.getsize(p) == 35434956


def test_library_exports_every_declared_symbol():
    lib = _lib.load()
    names = _lib.exported_symbols()
    assert len(names) >= 18
    for n in names:
        assert hasattr(lib, n), n
    assert lib.ay_abi_version() >= 1
    # no torch / C++ types in the header's signatures
    text = open(_lib.HEADER).read()
    assert "at::" not in text and "torch" not in text and "std::" not in text
    assert set(_lib._PROTOTYPES) | {"ay_last_error"} == set(names)


def test_layer_desc_struct_matches_header():
    text = open(_lib.HEADER).read()
    body = re.search(r"typedef struct ay_layer_desc \{(.*?)\} ay_layer_desc;", text, re.S).group(1)
    fields = re.findall(r"^\s*(int32_t|float)\s+(\w+)(?:\[(\d+)\])?;", body, re.M)
    assert [f[1] for f in fields] == [f[0] for f in _lib.LayerDesc._fields_]
    size = sum(4 * (int(n) if n else 1) for _, _, n in fields)
    assert ctypes.sizeof(_lib.LayerDesc) == size


def test_precision_enum_matches_header():
    """The host's activation-type constants are the header's (bf16 / fp16 tensor-core path, fp32 accuracy mode)."""
    text = open(_lib.HEADER).read()
    enum = re.search(r"enum \{ (AY_BF16 = \d+, AY_FP16 = \d+, AY_F32 = \d+) \};", text).group(1)
    vals = dict((k.strip(), int(v)) for k, v in (kv.split("=") for kv in enum.split(",")))
    assert vals == {"AY_BF16": _lib.AY_BF16, "AY_FP16": _lib.AY_FP16, "AY_F32": _lib.AY_F32}
    from amyloid_yolo_paper_b200.models import Darknet
    from amyloid_yolo_paper_b200.config import ensure_cfg
    assert Darknet(ensure_cfg("yolov3-tiny")).precision == "bf16"        # the tensor-core path is the default


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_product_fails_loudly_without_gpu(tmp_path):
    """No CPU fallback anywhere on the product path."""
    from amyloid_yolo_paper_b200.core import mergeDetections
    from amyloid_yolo_paper_b200.utils.utils import bbox_iou, non_max_suppression, rescale_boxes
    m = Darknet(write_mini_cfg(tmp_path), img_size=64)
    with pytest.raises(_lib.EngineUnavailable):
        m(torch.zeros(1, 3, 64, 64))
    with pytest.raises(_lib.EngineUnavailable):
        non_max_suppression(torch.zeros(1, 10, 7), 0.5, 0.4)
    with pytest.raises(_lib.EngineUnavailable):
        mergeDetections(torch.zeros(2, 7))
    with pytest.raises(_lib.EngineUnavailable):
        rescale_boxes(torch.zeros(2, 7), 416, (1536, 1536))
    with pytest.raises(_lib.EngineUnavailable):
        bbox_iou(torch.zeros(1, 4), torch.zeros(3, 4))
    handle = ctypes.c_void_p()
    assert _lib.load().ay_create(ctypes.byref(handle), 0) != 0
    assert b"no CPU fallback" in _lib.load().ay_last_error() or b"CUDA" in _lib.load().ay_last_error()
    with pytest.raises(NotImplementedError):
        m(torch.zeros(1, 3, 64, 64), targets=torch.zeros(1, 6))


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "amyloid_yolo_paper_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle", text, re.M), f
                assert "/root/reference" not in text, f


def test_detect_cli_flags_match_reference():
    from amyloid_yolo_paper_b200.detect import build_parser, pad_square
    p = build_parser()
    d = p.parse_args([])
    assert (d.conf_thres, d.nms_thres, d.batch_size, d.n_cpu, d.img_size) == (0.8, 0.4, 1, 0, 416)
    assert d.image_folder == "data/samples" and d.model_def == "config/yolov3.cfg"
    assert d.weights_path == "weights/yolov3.weights" and d.class_path == "data/coco.names"
    assert d.merge_boxes is None and d.checkpoint_model is None
    o = p.parse_args(["--merge_boxes", "True", "--img_size", "416", "--conf_thres", "0.8", "--nms_thres", "0.4"])
    assert o.merge_boxes == "True"                       # string, compared with "True" (detect.py:132)
    img = np.ones((4, 7, 3), np.uint8)
    sq = pad_square(img)
    assert sq.shape == (7, 7, 3) and sq[:1].sum() == 0 and sq[1:5].sum() == 4 * 7 * 3 and sq[5:].sum() == 0


def test_shard_ranges_cover_all_tiles():
    from amyloid_yolo_paper_b200.sweep import shard_range
    for n in (0, 1, 7, 8, 100000, 100003):
        for w in (1, 2, 4, 8):
            spans = [shard_range(n, r, w) for r in range(w)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(spans[i][1] == spans[i + 1][0] for i in range(w - 1))
            assert max(b - a for a, b in spans) - min(b - a for a, b in spans) <= 1


def test_loader_gather_equals_reference_transform_chain(tmp_path):
    """utils/datasets.py: gathering the sampled pixels on the uint8 image == PadSquare -> ToTensor -> Resize of the
    reference (utils/transforms.py:80-111) element for element, for square tiles and for both kinds of padding."""
    import torch.nn.functional as F
    from PIL import Image
    from amyloid_yolo_paper_b200.detect import pad_square
    from amyloid_yolo_paper_b200.utils.datasets import ImageFolder, TileLoader, gather_resized, nearest_index
    from oracle import restate as R
    rng = np.random.default_rng(0)
    for (h, w), S in (((1536, 1536), 416), ((300, 200), 64), ((201, 300), 96), ((50, 50), 64), ((640, 640), 320)):
        img = rng.integers(0, 256, size=(h, w, 3), dtype=np.uint8)
        sq = pad_square(img)
        x = torch.from_numpy(np.ascontiguousarray(sq)).permute(2, 0, 1).to(torch.float32).div(255)      # ToTensor
        want = F.interpolate(x.unsqueeze(0), size=S, mode="nearest").squeeze(0)                           # Resize
        got = torch.from_numpy(gather_resized(img, S)).permute(2, 0, 1).to(torch.float32).div(255)
        assert torch.equal(got, want), (h, w, S)
        assert np.array_equal(nearest_index(S, max(h, w)), R.nearest_src_index(S, max(h, w)))
    # the loader: sorted file order, batches of one original shape, pinned-or-not uint8 blocks
    shapes = [(40, 40), (40, 40), (40, 40), (30, 50), (40, 40)]
    imgs = []
    for i, (h, w) in enumerate(shapes):
        a = rng.integers(0, 256, size=(h, w, 3), dtype=np.uint8)
        imgs.append(a)
        Image.fromarray(a).save(str(tmp_path / ("t%02d.png" % i)))
    ds = ImageFolder(str(tmp_path), img_size=32)
    assert len(ds) == 5 and ds[3][2] == (30, 50) and ds[0][1].dtype == torch.uint8 and tuple(ds[0][1].shape) == (32, 32, 3)
    got = [(list(p), t.clone(), s) for p, t, s in TileLoader(str(tmp_path), img_size=32, batch_size=2, n_cpu=3)]
    assert [len(g[0]) for g in got] == [2, 1, 1, 1] and [g[2] for g in got] == [(40, 40), (40, 40), (30, 50), (40, 40)]
    flat = [t for g in got for t in g[1]]
    for a, t in zip(imgs, flat):
        assert np.array_equal(t.numpy(), gather_resized(a, 32))


def test_slide_tiling_host_logic(tmp_path):
    """grid / names / divide-and-conquer chunks against the oracle's branch-for-branch
    restatement, and save_and_tile's directory layout with white-padded edge tiles.  (Parity with pyvips itself is
    unpinned: it is not installed.)"""
    from PIL import Image
    from amyloid_yolo_paper_b200 import crop
    from oracle import restate as R
    for h, w in [(1536, 1536), (1537, 1535), (24576, 24576), (24577, 49152), (30000, 50000), (100, 73729), (73728, 24575)]:
        assert crop.divide_and_conquer_chunks(h, w) == R.divide_and_conquer_faithful(h, w), (h, w)
        rows, cols = crop.tile_grid(h, w)
        assert (rows - 1) * 1536 < h <= rows * 1536 and (cols - 1) * 1536 < w <= cols * 1536
    assert crop.tile_name(3, 7) == "0/3/7.jpg" and crop.chunk_basename(2, "s.svs") == "2_s.svs"
    rng = np.random.default_rng(0)
    slide = rng.integers(0, 200, (300, 450, 3), dtype=np.uint8)
    want = R.wsi_tiles_faithful(slide, 128)
    assert sorted(want) == [(r, c) for r in range(3) for c in range(4)]
    for (r, c), t in want.items():
        assert np.array_equal(crop.padded_tile(slide, r, c, 128), t)
    assert (want[(2, 3)][44:, :] == 255).all() and (want[(2, 3)][:, 66:] == 255).all()
    flat = np.full((300, 450, 3), 90, dtype=np.uint8)                       # flat colour survives JPEG: layout check
    files = crop.save_and_tile(flat, str(tmp_path), "slideA.svs", tile_size=128)
    assert len(files) == 12 and os.path.isfile(os.path.join(str(tmp_path), "slideA", "0", "2", "3.jpg"))
    edge = np.asarray(Image.open(os.path.join(str(tmp_path), "slideA", "0", "2", "3.jpg")))
    assert edge.shape == (128, 128, 3) and abs(int(edge[10, 10, 0]) - 90) <= 2 and int(edge[100, 100, 0]) >= 250
    assert crop.halve(np.full((5, 7, 3), 11, np.uint8)).shape == (2, 3, 3)
    with pytest.raises(ValueError):
        crop.detector_inputs(torch.zeros((10, 10, 3), dtype=torch.uint8), 0, 0, 1, 1)      # host tensor: no CPU path


def test_caa_host_side_fails_loudly_without_gpu():
    """The CAA filter's crop / network path has no CPU fallback either: host tensors are rejected, a missing GPU raises."""
    from amyloid_yolo_paper_b200 import caa
    tile = torch.zeros((600, 600, 3), dtype=torch.uint8)
    with pytest.raises(ValueError):
        caa.normalised_crops(tile, [(0, 256, 0, 256)], [0.5] * 3, [0.2] * 3)            # host tensor: no CPU path
    with pytest.raises(ValueError):
        caa.crop_window((10, 10, 5, 5), size=200)                                       # no 256-pixel window in a 200-pixel tile
    assert caa.crop_window((0, 0, 10, 10), 600) == (0, 256, 0, 256)
    assert caa.crop_window((590, 300, 10, 10), 600) == (344, 600, 177, 433)
    net = caa.Net()
    assert not net.training and net.precision == "fp32"
    if not torch.cuda.is_available():
        with pytest.raises(_lib.EngineUnavailable):
            net(torch.zeros(1, 3, 256, 256))
        with pytest.raises(_lib.EngineUnavailable):
            caa.filterDetectionsByCAAModel(tile, torch.zeros((1, 7)), ["Cored", "CAA"], model=net,
                                           norm={"mean": [0.5] * 3, "std": [0.2] * 3})
    empty = caa.filterDetectionsByCAAModel(tile, torch.zeros((0, 7)), ["Cored", "CAA"], model=net, norm={"mean": [0.5] * 3, "std": [0.2] * 3})
    assert empty.numel() == 0                                                           # nothing to classify: no engine needed


def test_caa_load_reference_model_from_a_pickled_module_or_state_dict(tmp_path):
    """core.py:432 pickles the whole nn.Module; load_reference_model takes its state_dict (or a saved state_dict)."""
    from amyloid_yolo_paper_b200 import caa
    from oracle.make_golden_caa import seeded_caa_state
    state = seeded_caa_state(2)
    src = caa.Net()
    src.load_state_dict(state)
    for obj, name in ((src, "module.pkl"), (state, "state.pkl")):
        path = str(tmp_path / name)
        torch.save(obj, path)
        net = caa.load_reference_model(path, precision="fp16")
        assert net.precision == "fp16" and not net.training
        got = net.state_dict()
        assert set(got) == set(state)
        assert all(torch.equal(got[k], state[k]) for k in state)
