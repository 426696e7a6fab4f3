"""lightcompress_b200 — B200-native (sm_100a) quantization hot path behind LLMC's quantizer API.

Drop-in for `llmc.compression.quantization`: IntegerQuantizer / FloatQuantizer and the RTN / GPTQ / Awq /
SmoothQuant ALGO_REGISTRY entries. Python host code -> ctypes -> C ABI (include/lcb200.h) -> hand-written
CUDA. No Triton, no multi-backend dispatch, no CPU fallback.
"""
from .quant import FloatQuantizer, IntegerQuantizer, weight_cast_to_bf16, weight_cast_to_fp8  # noqa: F401

__all__ = ['IntegerQuantizer', 'FloatQuantizer', 'weight_cast_to_bf16', 'weight_cast_to_fp8']
