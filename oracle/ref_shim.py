"""Import the UNMODIFIED reference quantizer (`audiocraft.quantization`).

TEST INFRASTRUCTURE ONLY.  Two places can hold it:
  * /root/reference (authoring container): used by tests/golden/make_golden.py and by CPU tests that skip
    when the reference is absent;
  * oracle/_ref/ (git-ignored copy of the four files of the path, staged by oracle/fetch_ref.py and shipped
    to the GPU box): what `bench.py --impl reference` and bench.py's cpu_baseline leg time on the host cores.
    /root/reference itself does not exist on the GPU box.

Two obstacles, neither needs a change to the reference:
  * `audiocraft/__init__.py` imports data/models/modules, which need packages that are not
    installed -> register an empty `audiocraft` package whose __path__ points at the reference so
    that `audiocraft.quantization` imports without running that __init__.
  * `core_vq.py:10` imports `flashy` (not installed) -> a stub whose distrib.broadcast_tensors /
    average_tensors do nothing (single process: nothing to synchronise).
"""
from __future__ import annotations

import os
import sys
import types

_STAGED = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")


def _has(root: str) -> bool:
    return os.path.isfile(os.path.join(root, "audiocraft", "quantization", "core_vq.py"))


def _pick_root() -> str:
    env = os.environ.get("RVQ_REFERENCE_ROOT")
    if env:
        return env
    if _has("/root/reference"):
        return "/root/reference"
    if _has(_STAGED):
        from . import fetch_ref
        if not fetch_ref.verify():
            raise RuntimeError("oracle/_ref does not match its manifest: run oracle/fetch_ref.py again")
        return _STAGED
    return "/root/reference"


REFERENCE_ROOT = _pick_root()


def available() -> bool:
    return _has(REFERENCE_ROOT)


def load():
    """Returns the reference's `audiocraft.quantization` module."""
    if not available():
        raise RuntimeError(f"reference not found under {REFERENCE_ROOT}")
    if "flashy" not in sys.modules:
        flashy = types.ModuleType("flashy")
        distrib = types.ModuleType("flashy.distrib")
        distrib.broadcast_tensors = lambda tensors, src=0: None
        distrib.average_tensors = lambda tensors: None
        distrib.is_distributed = lambda: False
        distrib.world_size = lambda: 1
        distrib.rank = lambda: 0
        flashy.distrib = distrib
        sys.modules["flashy"] = flashy
        sys.modules["flashy.distrib"] = distrib
    if "audiocraft" not in sys.modules or not hasattr(sys.modules["audiocraft"], "__path__"):
        pkg = types.ModuleType("audiocraft")
        pkg.__path__ = [os.path.join(REFERENCE_ROOT, "audiocraft")]
        sys.modules["audiocraft"] = pkg
    import importlib
    return importlib.import_module("audiocraft.quantization")
