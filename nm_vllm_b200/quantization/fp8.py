"""W8A8 FP8 (per-tensor) linear and the fp8 KV-cache scale: mirror of vllm/model_executor/layers/quantization/fp8.py
(`Fp8Config` :35-88, `Fp8LinearMethod` :92-379, `Fp8KVCacheMethod` :563-598, `per_tensor_quantize/dequantize` :601-613).
On sm_100a the FP8 tensor path always exists, so the reference's Marlin weight-only fallback for pre-Ada GPUs (:108-111,
:208-247) and the torch._scaled_mm branch (:358-377) have no counterpart: `apply` is scaled_fp8_quant (static or
dynamic) followed by cutlass_scaled_mm, with the bias fused into the GEMM epilogue."""
from typing import Any, Dict, List, Optional, Union

import torch
from torch.nn import Module
from torch.nn.parameter import Parameter

from .. import _custom_ops as ops
from .base import LinearBase, LinearMethodBase, QuantizationConfig, QuantizeMethodBase, set_weight_attrs

ACTIVATION_SCHEMES = ["static", "dynamic"]


def cutlass_fp8_supported() -> bool:
    return True   # tcgen05 kind::f8f6f4 (csrc/scaled_mm.cu); ops.cutlass_scaled_mm_supports_fp8(100) says the same


class Fp8Config(QuantizationConfig):
    def __init__(self, is_checkpoint_fp8_serialized: bool = False, activation_scheme: str = "dynamic") -> None:
        if activation_scheme not in ACTIVATION_SCHEMES:
            raise ValueError(f"Unsupported activation scheme {activation_scheme}")
        self.is_checkpoint_fp8_serialized = is_checkpoint_fp8_serialized
        self.activation_scheme = activation_scheme

    @classmethod
    def get_name(cls) -> str:
        return "fp8"

    @classmethod
    def get_supported_act_dtypes(cls) -> List[torch.dtype]:
        return [torch.bfloat16, torch.half]

    @classmethod
    def get_min_capability(cls) -> int:
        return 80

    @classmethod
    def get_config_filenames(cls) -> List[str]:
        return []

    @classmethod
    def from_config(cls, config: Dict[str, Any]) -> "Fp8Config":
        return cls(is_checkpoint_fp8_serialized="fp8" in cls.get_from_keys(config, ["quant_method"]),
                   activation_scheme=cls.get_from_keys(config, ["activation_scheme"]))

    def get_quant_method(self, layer: torch.nn.Module) -> Optional[QuantizeMethodBase]:
        from ..attention.layer import Attention
        if isinstance(layer, LinearBase):
            return Fp8LinearMethod(self)
        if isinstance(layer, Attention):
            return Fp8KVCacheMethod(self)
        return None


def per_tensor_quantize(tensor: torch.Tensor, inv_scale: Union[float, torch.Tensor]) -> torch.Tensor:
    finfo = torch.finfo(torch.float8_e4m3fn)
    return (tensor / inv_scale).clamp(min=finfo.min, max=finfo.max).to(torch.float8_e4m3fn)


def per_tensor_dequantize(tensor: torch.Tensor, inv_scale: Union[float, torch.Tensor]) -> torch.Tensor:
    return tensor.to(torch.float16) * inv_scale


class Fp8LinearMethod(LinearMethodBase):
    def __init__(self, quant_config: Fp8Config):
        self.quant_config = quant_config
        self.cutlass_fp8_supported = cutlass_fp8_supported()
        self.use_marlin = False

    def _create_scale_param(self, scale_name: str, layer: Module, output_partition_sizes: List[int], **extra_weight_attrs) -> None:
        # one scale per LOGICAL weight of a fused layer (q / k / v, gate / up); "min" marks a slot the checkpoint did not fill
        scale = Parameter(torch.empty(len(output_partition_sizes), dtype=torch.float32), requires_grad=False)
        scale[:] = torch.finfo(torch.float8_e4m3fn).min
        layer.register_parameter(scale_name, scale)
        set_weight_attrs(scale, {**extra_weight_attrs, "needs_scalar_to_array": True})

    def create_weights(self, layer: Module, input_size_per_partition: int, output_partition_sizes: List[int], input_size: int,
                       output_size: int, params_dtype: torch.dtype, **extra_weight_attrs):
        del input_size, output_size
        n_part = sum(output_partition_sizes)
        layer.process_after_load = True
        layer.logical_widths = output_partition_sizes
        layer.input_size_per_partition = input_size_per_partition
        layer.output_size_per_partition = n_part
        layer.orig_dtype = params_dtype
        serialized = self.quant_config.is_checkpoint_fp8_serialized
        weight = Parameter(torch.empty(n_part, input_size_per_partition, dtype=torch.float8_e4m3fn if serialized else params_dtype),
                           requires_grad=False)
        layer.register_parameter("weight", weight)
        set_weight_attrs(weight, {**extra_weight_attrs, "input_dim": 1, "output_dim": 0})
        if serialized:
            self._create_scale_param("weight_scale", layer, output_partition_sizes, **extra_weight_attrs)
            if self.quant_config.activation_scheme == "static":
                self._create_scale_param("input_scale", layer, output_partition_sizes, **extra_weight_attrs)

    def process_weights_after_loading(self, layer: Module) -> None:
        if not getattr(layer, "process_after_load", False):
            return
        if not self.quant_config.is_checkpoint_fp8_serialized:
            # 16-bit checkpoint: quantise the whole (fused) weight with one dynamic per-tensor scale, our own kernel
            qweight, weight_scale = ops.scaled_fp8_quant(layer.weight, scale=None)
            layer.weight = Parameter(qweight.t(), requires_grad=False)      # [K, N] column-major view for cutlass_scaled_mm
            layer.weight_scale = Parameter(weight_scale, requires_grad=False)
            layer.logical_widths = None
            layer.input_scale = None
            return
        # fp8 checkpoint: the logical weights of a fused layer were quantised separately -> requantise to ONE scale (the
        # largest), unless the checkpoint stored the layer fused (then only slot 0 was filled, fp8.py:286-296)
        max_w_scale = layer.weight_scale.max()
        if layer.weight_scale[-1] > torch.finfo(torch.float8_e4m3fn).min:
            start = 0
            for idx, width in enumerate(layer.logical_widths):
                dq = per_tensor_dequantize(layer.weight[start:start + width, :], layer.weight_scale[idx])
                layer.weight[start:start + width, :] = per_tensor_quantize(dq, max_w_scale)
                start += width
        layer.weight_scale = Parameter(max_w_scale, requires_grad=False)
        layer.weight = Parameter(layer.weight.t(), requires_grad=False)
        if self.quant_config.activation_scheme == "dynamic":
            layer.input_scale = None
        elif self.quant_config.activation_scheme == "static":
            layer.input_scale = Parameter(layer.input_scale.max(), requires_grad=False)
        else:
            raise ValueError(f"Unknown scheme {self.quant_config.activation_scheme}")

    def apply(self, layer: Module, x: torch.Tensor, bias: Optional[torch.Tensor] = None) -> torch.Tensor:
        x2 = x.reshape(-1, x.shape[-1])
        qinput, x_scale = ops.scaled_fp8_quant(x2, layer.input_scale)   # input_scale None -> dynamic per-tensor
        out = ops.cutlass_scaled_mm(qinput, layer.weight, x_scale, layer.weight_scale, x.dtype, bias)
        return out.reshape(x.shape[:-1] + (out.shape[-1],))


class Fp8KVCacheMethod(QuantizeMethodBase):
    """Loads the per-tensor kv_scale of an attention layer from an fp8 checkpoint (fp8.py:563-598): a scalar parameter
    `kv_scale` (default 1.0) that becomes the python float `layer._kv_scale` the attention ops are called with."""

    def __init__(self, quant_config: Fp8Config):
        self.quant_config = quant_config

    def create_weights(self, layer: Module):
        layer.kv_scale = Parameter(torch.tensor(1.0), requires_grad=False)

    def apply(self, layer: Module) -> torch.Tensor:
        raise RuntimeError("Fp8KVCacheMethod.apply should not be called.")

    def process_weights_after_loading(self, layer: Module) -> None:
        if layer.kv_cache_dtype != "auto":   # with an "auto" cache the scale stays 1.0 whatever the checkpoint holds
            kv_scale = layer.kv_scale.to("cpu").tolist()
            if not isinstance(kv_scale, float):
                raise ValueError("Only support per-tensor scaling factor for fp8 KV cache")
            layer._kv_scale = kv_scale
        del layer.kv_scale
