"""QuantizationConfig / LinearMethodBase mirrors for the formats on the hot path."""
from .base import LinearBase, LinearMethodBase, QuantizationConfig, QuantizeMethodBase, set_weight_attrs
from .compressed_tensors import (CompressedTensorsConfig, CompressedTensorsLinearMethod, CompressedTensorsW4A16Sparse24,
                                 CompressedTensorsW8A8, CompressedTensorsWNA16)
from .fp8 import Fp8Config, Fp8KVCacheMethod, Fp8LinearMethod
from .gptq_marlin import GPTQMarlinConfig, GPTQMarlinLinearMethod, GPTQMarlinState
from .layouts import get_scale_perms, marlin_permute_scales
from . import sparse24
from .marlin import GPTQMarlin24Config, GPTQMarlin24LinearMethod, MarlinConfig, MarlinLinearMethod

QUANTIZATION_METHODS = {
    "fp8": Fp8Config,
    "marlin": MarlinConfig,
    "gptq_marlin_24": GPTQMarlin24Config,
    "gptq_marlin": GPTQMarlinConfig,
    "sparseml": CompressedTensorsConfig,
    "compressed-tensors": CompressedTensorsConfig,
}


def get_quantization_config(quantization: str):
    if quantization not in QUANTIZATION_METHODS:
        raise ValueError(f"Invalid quantization method: {quantization}")
    return QUANTIZATION_METHODS[quantization]
