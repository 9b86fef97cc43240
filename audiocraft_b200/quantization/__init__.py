"""Mirror of the reference package surface (audiocraft/quantization/__init__.py:8-9)."""
# flake8: noqa
from .vq import ResidualVectorQuantizer
from .base import BaseQuantizer, DummyQuantizer, QuantizedResult
