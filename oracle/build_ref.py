"""TEST / MEASUREMENT INFRASTRUCTURE ONLY — packs the reference's own hot-path modules for the GPU box.

The reference is pure Python: there is nothing to compile.  So that ``bench.py --impl reference`` can time the
UNMODIFIED reference (``cpu_baseline.kind = "reference"``) on a box where ``/root/reference`` does not exist, this
script zips the five source files the path imports — straight from where they lie under ``/root/reference`` — into
``oracle/_ref/reference_py.zip`` (git-ignored: the sources never enter this repository's history; the directory is
not gpurun-ignored, so the archive travels to the GPU box like the built ``.so``).  ``oracle/ref_import.py`` imports
from the checkout when it is mounted and from this archive otherwise.

    python -m oracle.build_ref          (also run by __graft_entry__.build() when the checkout is present)
"""
import os
import zipfile

from . import ref_import

FILES = ["models.py", "core.py", "utils/__init__.py", "utils/utils.py", "utils/parse_config.py", "LICENSE"]
ARCHIVE = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref", "reference_py.zip")


def build():
    root = ref_import.REFERENCE_ROOT
    if not os.path.isfile(os.path.join(root, "models.py")):
        return None
    os.makedirs(os.path.dirname(ARCHIVE), exist_ok=True)
    tmp = ARCHIVE + ".tmp%d" % os.getpid()
    with zipfile.ZipFile(tmp, "w", zipfile.ZIP_DEFLATED) as z:
        for rel in FILES:
            src = os.path.join(root, rel)
            if os.path.isfile(src):
                z.write(src, rel)
            elif rel.endswith("__init__.py"):
                z.writestr(rel, "")
    os.replace(tmp, ARCHIVE)
    return ARCHIVE


if __name__ == "__main__":
    print(build())
