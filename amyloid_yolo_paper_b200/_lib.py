"""ctypes binding of ``libamyloid_yolo_b200.so`` (C ABI: ``include/amyloid_yolo.h``).

The library is built in-tree by :func:`build` (``nvcc -gencode
arch=compute_100a,code=sm_100a``) and loaded with ctypes; there is no pybind /
torch extension in between and no fallback: if the library is missing or no
sm_100 GPU is present the product raises :class:`EngineUnavailable`.
"""
import ctypes
import os
import subprocess
import threading

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB_PATH = os.path.join(HERE, "libamyloid_yolo_b200.so")
HEADER = os.path.join(os.path.dirname(HERE), "include", "amyloid_yolo.h")
SOURCES = ["engine.cu", "conv_gemm_sm100.cu", "first_conv_ts.cu", "aux_kernels.cu", "fp32_path.cu", "nms.cu", "merge.cu"]

AY_CONV, AY_MAXPOOL, AY_UPSAMPLE, AY_ROUTE, AY_SHORTCUT, AY_YOLO, AY_LINEAR = range(7)
AY_IN_F32_NCHW, AY_IN_U8_HWC = 0, 1
AY_BF16, AY_FP16, AY_F32 = 0, 1, 2


class EngineUnavailable(RuntimeError):
    pass


class EngineError(RuntimeError):
    def __init__(self, code, message):
        super().__init__("amyloid_yolo_b200 error %d: %s" % (code, message))
        self.code = code


class LayerDesc(ctypes.Structure):
    _fields_ = [("type", ctypes.c_int32), ("filters", ctypes.c_int32), ("size", ctypes.c_int32),
                ("stride", ctypes.c_int32), ("leaky", ctypes.c_int32), ("n_src", ctypes.c_int32),
                ("src", ctypes.c_int32 * 4), ("n_anchors", ctypes.c_int32), ("classes", ctypes.c_int32),
                ("anchors", ctypes.c_float * 16)]


def nvcc_command(out_path=LIB_PATH):
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    return [nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
            "-Xcompiler", "-fPIC", "-shared", "-o", out_path] + [os.path.join(CSRC, s) for s in SOURCES]


def needs_build():
    if not os.path.exists(LIB_PATH):
        return True
    built = os.path.getmtime(LIB_PATH)
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + [HEADER]
    return any(os.path.getmtime(d) > built for d in deps)


def build(force=False, verbose=False):
    """Compile the CUDA sources for sm_100a into the in-tree shared library."""
    if not force and not needs_build():
        return LIB_PATH
    tmp = LIB_PATH + ".tmp%d" % os.getpid()
    cmd = nvcc_command(tmp)
    if verbose:
        print(" ".join(cmd))
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + res.stdout + res.stderr)
    os.replace(tmp, LIB_PATH)
    return LIB_PATH


_lock = threading.Lock()
_lib = None

_PROTOTYPES = {
    "ay_create": [ctypes.POINTER(ctypes.c_void_p), ctypes.c_int],
    "ay_destroy": [ctypes.c_void_p],
    "ay_abi_version": [],
    "ay_net_build": [ctypes.c_void_p, ctypes.POINTER(LayerDesc), ctypes.c_int, ctypes.c_int, ctypes.c_int,
                     ctypes.c_int, ctypes.c_int],
    "ay_net_set_conv": [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p],
    "ay_net_finalize": [ctypes.c_void_p],
    "ay_net_output_shape": [ctypes.c_void_p, ctypes.POINTER(ctypes.c_int), ctypes.POINTER(ctypes.c_int)],
    "ay_forward": [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                   ctypes.c_void_p, ctypes.c_void_p],
    "ay_net_layer_output": [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p],
    "ay_net_layer_shape": [ctypes.c_void_p, ctypes.c_int, ctypes.POINTER(ctypes.c_int),
                           ctypes.POINTER(ctypes.c_int), ctypes.POINTER(ctypes.c_int)],
    "ay_net_layer_impl": [ctypes.c_void_p, ctypes.c_int],
    "ay_net_launch_count": [ctypes.c_void_p],
    "ay_nms": [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_float,
               ctypes.c_float, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p],
    "ay_forward_candidates": [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                              ctypes.c_void_p, ctypes.c_float, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p],
    "ay_nms_candidates": [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int,
                          ctypes.c_int, ctypes.c_int, ctypes.c_float, ctypes.c_float, ctypes.c_void_p, ctypes.c_void_p,
                          ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p],
    "ay_nms_timed": [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_float,
                     ctypes.c_float, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p,
                     ctypes.POINTER(ctypes.c_float)],
    "ay_nms_static_capacity": [],
    "ay_nms_static": [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_float, ctypes.c_float,
                      ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p],
    "ay_rescale_merge": [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int,
                         ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_void_p,
                         ctypes.c_void_p, ctypes.c_void_p],
    "ay_crop_normalize": [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_longlong, ctypes.c_void_p,
                          ctypes.c_int, ctypes.c_int, ctypes.POINTER(ctypes.c_float), ctypes.POINTER(ctypes.c_float),
                          ctypes.c_void_p, ctypes.c_void_p],
    "ay_wsi_gather": [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_longlong, ctypes.c_int,
                      ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_int,
                      ctypes.c_void_p],
    "ay_bbox_iou": [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int,
                    ctypes.c_void_p, ctypes.c_void_p],
    "ay_check": [ctypes.c_void_p, ctypes.c_void_p],
    "ay_net_use_reference_conv": [ctypes.c_void_p, ctypes.c_int],
    "ay_forward_profile": [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                           ctypes.c_void_p, ctypes.c_void_p, ctypes.POINTER(ctypes.c_float), ctypes.c_int],
    "ay_debug_set": [ctypes.c_void_p, ctypes.c_int],
    "ay_prof_words": [],
    "ay_forward_stamps": [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                          ctypes.c_void_p, ctypes.c_void_p, ctypes.POINTER(ctypes.c_ulonglong), ctypes.c_int],
    "ay_debug_counters": [ctypes.c_void_p, ctypes.POINTER(ctypes.c_ulonglong), ctypes.c_int],
}


def load():
    """dlopen the library (no GPU needed for this) and set prototypes."""
    global _lib
    with _lock:
        if _lib is not None:
            return _lib
        if not os.path.exists(LIB_PATH):
            raise EngineUnavailable(
                "CUDA library %s is missing; run `python -c 'import __graft_entry__ as g; g.build()'` "
                "(there is no CPU fallback)" % LIB_PATH)
        lib = ctypes.CDLL(LIB_PATH)
        for name, args in _PROTOTYPES.items():
            fn = getattr(lib, name)
            fn.argtypes = args
            fn.restype = ctypes.c_int
        lib.ay_nms_static_capacity.restype = ctypes.c_size_t
        lib.ay_last_error.argtypes = []
        lib.ay_last_error.restype = ctypes.c_char_p
        _lib = lib
        return lib


def check(code):
    if code != 0:
        raise EngineError(code, load().ay_last_error().decode("utf-8", "replace"))


def exported_symbols():
    """Function names declared in include/amyloid_yolo.h (for the load test)."""
    import re
    text = open(HEADER).read()
    return sorted(set(re.findall(r"\b(ay_[a-z0-9_]+)\s*\(", text)))


_ctx_lock = threading.Lock()
_contexts = {}


def context(device_index):
    """One ay_ctx per GPU, created on first use."""
    lib = load()
    with _ctx_lock:
        if device_index not in _contexts:
            handle = ctypes.c_void_p()
            code = lib.ay_create(ctypes.byref(handle), int(device_index))
            if code != 0:
                raise EngineUnavailable(lib.ay_last_error().decode("utf-8", "replace"))
            _contexts[device_index] = handle
        return _contexts[device_index]
