"""audiocraft_b200 -- B200-native EnCodec RVQ tokenizer path (drop-in for audiocraft.quantization).

Only the hot path named in BASELINE.json lives here: `quantization.ResidualVectorQuantizer`
(encode / decode / forward -> QuantizedResult) on hand-written sm_100a kernels behind the C ABI of
include/rvq_b200.h, plus the frame-sharded multi-GPU driver in `sharded`.
"""
__version__ = "0.1.0"

from .quantization import (  # noqa: F401
    BaseQuantizer, DummyQuantizer, QuantizedResult, ResidualVectorQuantizer,
)
