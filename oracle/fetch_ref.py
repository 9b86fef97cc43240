"""Stage the UNMODIFIED reference quantizer for the GPU box (TEST INFRASTRUCTURE ONLY).

    python oracle/fetch_ref.py          # authoring container: /root/reference -> oracle/_ref/

Copies the four files of the path, byte for byte, from where they lie under the reference,

    /root/reference/audiocraft/quantization/{__init__,base,vq,core_vq}.py

into `oracle/_ref/audiocraft/quantization/`.  `oracle/_ref/` is git-ignored (no reference source enters
the history) but not gpurun-ignored, so the copy travels to the GPU box with the snapshot, where
`bench.py --impl reference` and `bench.py`'s cpu_baseline leg time it on the host cores through
`oracle/ref_shim.py` (kind "reference").  A sha256 manifest is written beside the copy; the shim refuses a
copy whose files do not match it.  Nothing in the product imports any of this.
"""
from __future__ import annotations

import hashlib
import json
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
SRC_ROOT = os.environ.get("RVQ_REFERENCE_SRC", "/root/reference")
DST_ROOT = os.path.join(HERE, "_ref")
FILES = ["__init__.py", "base.py", "vq.py", "core_vq.py"]
REL = os.path.join("audiocraft", "quantization")


def sha256(path: str) -> str:
    with open(path, "rb") as f:
        return hashlib.sha256(f.read()).hexdigest()


def fetch(quiet: bool = False) -> bool:
    """Returns True when oracle/_ref holds a verified copy afterwards."""
    src_dir = os.path.join(SRC_ROOT, REL)
    dst_dir = os.path.join(DST_ROOT, REL)
    if not os.path.isdir(src_dir):
        return verify()
    os.makedirs(dst_dir, exist_ok=True)
    manifest = {}
    for name in FILES:
        shutil.copyfile(os.path.join(src_dir, name), os.path.join(dst_dir, name))
        manifest[name] = sha256(os.path.join(dst_dir, name))
        assert manifest[name] == sha256(os.path.join(src_dir, name))
    with open(os.path.join(DST_ROOT, "MANIFEST.json"), "w") as f:
        json.dump({"source": src_dir, "sha256": manifest}, f, indent=1)
    if not quiet:
        print(f"copied {len(FILES)} files from {src_dir} to {dst_dir}")
    return True


def verify() -> bool:
    path = os.path.join(DST_ROOT, "MANIFEST.json")
    if not os.path.isfile(path):
        return False
    with open(path) as f:
        manifest = json.load(f)["sha256"]
    dst_dir = os.path.join(DST_ROOT, REL)
    return all(os.path.isfile(os.path.join(dst_dir, n)) and sha256(os.path.join(dst_dir, n)) == h
               for n, h in manifest.items()) and set(manifest) == set(FILES)


if __name__ == "__main__":
    sys.exit(0 if fetch() else 1)
