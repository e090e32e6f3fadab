"""TEST INFRASTRUCTURE ONLY — loader for the *unmodified* reference sources.

Only usable in the build container, where the reference checkout is mounted
read-only at /root/reference.  It is imported by ``oracle/make_golden.py`` (to
generate ``tests/golden/*``) and by the container-only tests that pin the
restatement in ``oracle/restate.py`` against the real thing.  Nothing under
``amyloid_yolo_paper_b200/`` may import this module.

The reference's hot-path modules import plotting packages that are absent here
(``models.py:12-13`` and ``utils/utils.py:10-11`` -> matplotlib; ``core.py:10``
-> skimage).  None of them is touched on the detection path, so inert stand-in
modules are registered before the import.  ``detect.py`` / ``validation.py`` /
``prospective.py`` must NOT be imported: they delete ``output/`` and run whole
analyses at import time (``detect.py:27``, ``validation.py:407-417``).
"""
import importlib
import os
import sys
import types

REFERENCE_ROOT = os.environ.get("AMYLOID_REFERENCE_ROOT", "/root/reference")

_ABSENT = [
    "matplotlib", "matplotlib.pyplot", "matplotlib.patches", "matplotlib.ticker",
    "matplotlib.colors", "matplotlib.lines", "matplotlib.cm",
    "skimage", "skimage.io", "skimage.transform", "skimage.measure", "skimage.draw",
    "imgaug", "imgaug.augmenters", "imgaug.augmentables", "imgaug.augmentables.bbs",
    "terminaltables", "statsmodels", "statsmodels.api", "pyvips",
]


class _Inert(types.ModuleType):
    """Module whose every public attribute is a do-nothing callable."""

    def __getattr__(self, name):
        if name.startswith("__") and name.endswith("__"):
            raise AttributeError(name)
        return _Inert(self.__name__ + "." + name)

    def __call__(self, *a, **k):
        return _Inert(self.__name__ + "()")


ARCHIVE = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref", "reference_py.zip")


def checkout_available():
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "models.py"))


def available():
    """The mounted checkout (build container) or the archive ``oracle/build_ref.py`` packed from it (GPU box)."""
    return checkout_available() or os.path.isfile(ARCHIVE)


def source():
    return "checkout" if checkout_available() else ("archive" if os.path.isfile(ARCHIVE) else None)


_cache = {}


def load():
    """-> (models, utils.utils, core) modules of the reference, unmodified."""
    if "mods" in _cache:
        return _cache["mods"]
    if not available():
        raise RuntimeError("reference sources not found (%s, %s)" % (REFERENCE_ROOT, ARCHIVE))
    root = REFERENCE_ROOT if checkout_available() else ARCHIVE        # zipimport handles the archive
    for name in _ABSENT:
        try:
            importlib.import_module(name)
        except Exception:
            mod = _Inert(name)
            mod.__path__ = []  # behave like a package for "import a.b"
            sys.modules[name] = mod
    # the reference uses top-level names ("utils", "models", "core"); make sure
    # ours do not shadow them and vice versa.
    saved = {k: sys.modules.pop(k) for k in list(sys.modules)
             if k in ("utils", "models", "core") or k.startswith("utils.")}
    sys.path.insert(0, root)
    cwd = os.getcwd()
    try:
        if os.path.isdir(root):
            os.chdir(root)
        models = importlib.import_module("models")
        uu = importlib.import_module("utils.utils")
        core = importlib.import_module("core")
    finally:
        os.chdir(cwd)
        sys.path.remove(root)
    ref = {k: sys.modules.pop(k) for k in list(sys.modules)
           if k in ("utils", "models", "core") or k.startswith("utils.")}
    _cache["ref_modules"] = ref
    sys.modules.update(saved)
    _cache["mods"] = (models, uu, core)
    return _cache["mods"]
