"""compressed-tensors checkpoints on the hot path: mirror of
vllm/model_executor/layers/quantization/compressed_tensors/compressed_tensors.py (`CompressedTensorsConfig` :19-188,
`CompressedTensorsLinearMethod` :189-234), utils.py (format / strategy enums, `QuantizationArgs`,
`find_first_name_or_class_match`) and the three schemes the path uses: `CompressedTensorsWNA16`
(schemes/compressed_tensors_wNa16.py:18-175, pack-quantized -> Marlin W4A16 / W8A16), `CompressedTensorsW4A16Sparse24`
(_w4a16_24.py:17-138, marlin-24 -> Sparse-Marlin) and `CompressedTensorsW8A8` (_w8a8.py:14-109, int-quantized ->
cutlass_scaled_mm INT8, static per-tensor or dynamic per-token activations)."""
import re
from dataclasses import dataclass
from enum import Enum
from typing import Any, Callable, Dict, Iterable, List, Optional

import torch
from torch.nn import Module, Parameter

from .. import _custom_ops as ops
from .base import LinearBase, LinearMethodBase, QuantizationConfig, set_weight_attrs
from .gptq_marlin import GPTQMarlinState, replace_tensor
from .layouts import GPTQ_MARLIN_MAX_PARALLEL, GPTQ_MARLIN_MIN_THREAD_N, marlin_permute_scales
from .marlin import GPTQ_MARLIN_24_MAX_PARALLEL, GPTQ_MARLIN_24_MIN_THREAD_N

WNA16_SUPPORTED_BITS = [4, 8]
W4A16SPARSE24_SUPPORTED_BITS = [4]


class CompressionFormat(Enum):
    dense = "dense"
    sparse_bitmask = "sparse-bitmask"
    int_quantized = "int-quantized"
    pack_quantized = "pack-quantized"
    marlin_24 = "marlin-24"


class QuantizationType(str, Enum):
    INT = "int"
    FLOAT = "float"


class QuantizationStrategy(str, Enum):
    TENSOR = "tensor"
    CHANNEL = "channel"
    GROUP = "group"
    BLOCK = "block"
    TOKEN = "token"


@dataclass
class QuantizationArgs:
    """One `weights` / `input_activations` entry of a config group (utils.py:35-67; a plain dataclass, no pydantic)."""
    num_bits: int = 8
    type: str = "int"
    symmetric: bool = True
    group_size: Optional[int] = None
    strategy: Optional[str] = None
    block_structure: Optional[str] = None
    dynamic: bool = False
    observer: str = "minmax"

    @classmethod
    def parse_obj(cls, obj: Optional[Dict[str, Any]]) -> "QuantizationArgs":
        if obj is None:
            raise ValueError("no quantization arguments")
        known = {k: v for k, v in obj.items() if k in cls.__dataclass_fields__}
        a = cls(**known)
        if isinstance(a.strategy, Enum):
            a.strategy = a.strategy.value
        return a


def _match(value: str, targets: Iterable[str], check_contains: bool) -> Optional[str]:
    for t in targets:
        if t.startswith("re:"):
            if re.match(t[3:], value):
                return t
        elif t == value or (check_contains and t.lower() in value.lower()):
            return t
    return None


def find_first_name_or_class_match(name: str, module: Module, targets: Iterable[str], check_contains: bool = False) -> Optional[str]:
    """The first target that matches the layer's name or its class name ("Linear" matches ColumnParallelLinear when
    check_contains, utils.py:84-122)."""
    return _match(name, targets, False) or _match(module.__class__.__name__, targets, check_contains)


class CompressedTensorsScheme:
    def create_weights(self, *args, **kwargs):
        raise NotImplementedError

    def apply_weights(self, layer: Module, x: torch.Tensor):
        raise NotImplementedError

    def process_weights_after_loading(self, layer: Module) -> None:
        raise NotImplementedError


class CompressedTensorsWNA16(CompressedTensorsScheme):
    """weight_packed [N, K / pf] int32 + weight_scale [N, K / g] (+ weight_shape): transposed GPTQ packing; the first
    apply_weights turns it into the Marlin layout exactly like GPTQMarlinLinearMethod does."""

    def __init__(self, strategy: str, num_bits: int, group_size: Optional[int] = None):
        if strategy == "group" and group_size is None:
            raise ValueError("group_size must be given when using strategy group")
        self.num_bits, self.strategy, self.group_size = num_bits, strategy, group_size

    def process_weights_after_loading(self, layer: Module) -> None:
        pass

    def create_weights(self, layer: Module, input_size: int, output_partition_sizes: List[int], input_size_per_partition: int,
                       params_dtype: torch.dtype, weight_loader: Callable, **kwargs):
        pf, n_part = 32 // self.num_bits, sum(output_partition_sizes)
        group = self.group_size if self.group_size is not None else input_size
        scale_dim, n_groups = None, input_size // group
        if input_size != input_size_per_partition and self.group_size is not None:
            scale_dim, n_groups = 1, input_size_per_partition // group     # row-parallel: scales shard with K
        weight = Parameter(torch.empty(n_part, input_size_per_partition // pf, dtype=torch.int32), requires_grad=False)
        set_weight_attrs(weight, {"input_dim": 1, "output_dim": 0, "packed_dim": 1, "pack_factor": pf, "weight_loader": weight_loader})
        layer.register_parameter("weight_packed", weight)
        weight_scale = Parameter(torch.empty(n_part, n_groups, dtype=params_dtype), requires_grad=False)
        set_weight_attrs(weight_scale, {"weight_loader": weight_loader, "input_dim": scale_dim, "output_dim": 0})
        layer.register_parameter("weight_scale", weight_scale)
        weight_shape = Parameter(torch.empty(2, dtype=torch.int64), requires_grad=False)   # original [N, K] before packing
        layer.register_parameter("weight_shape", weight_shape)
        set_weight_attrs(weight_shape, {"weight_loader": weight_loader, "ignore_warning": True})
        layer.input_size_per_partition, layer.output_size_per_partition = input_size_per_partition, n_part
        layer.input_size = input_size
        layer.marlin_state = GPTQMarlinState.REPACK
        layer.is_k_full = True
        layer.group_size = group
        layer.workspace = torch.zeros((n_part // GPTQ_MARLIN_MIN_THREAD_N) * GPTQ_MARLIN_MAX_PARALLEL, dtype=torch.int, requires_grad=False)

    def apply_weights(self, layer: Module, x: torch.Tensor):
        x2 = x.reshape(-1, x.shape[-1])
        n_part, k_part = layer.output_size_per_partition, layer.input_size_per_partition
        if layer.marlin_state == GPTQMarlinState.REPACK:
            layer.marlin_state = GPTQMarlinState.READY
            dev = layer.weight_packed.device
            layer.g_idx = Parameter(torch.empty(0, dtype=torch.int, device=dev), requires_grad=False)
            layer.g_idx_sort_indices = Parameter(torch.empty(0, dtype=torch.int, device=dev), requires_grad=False)
            replace_tensor(layer, "weight_packed", ops.gptq_marlin_repack(layer.weight_packed.t().contiguous(), layer.g_idx_sort_indices,
                                                                           k_part, n_part, self.num_bits))
            replace_tensor(layer, "weight_scale", marlin_permute_scales(layer.weight_scale.squeeze().t().contiguous(), k_part, n_part,
                                                                         layer.group_size, self.num_bits))
            layer.workspace = layer.workspace.to(dev)
        out = ops.gptq_marlin_gemm(x2, layer.weight_packed, layer.weight_scale, layer.g_idx, layer.g_idx_sort_indices, layer.workspace,
                                   self.num_bits, x2.shape[0], n_part, k_part, layer.is_k_full)
        return out.reshape(x.shape[:-1] + (n_part,))


class CompressedTensorsW4A16Sparse24(CompressedTensorsScheme):
    """weight_packed / scale_packed / meta (+ weight_shape) already in Sparse-Marlin layout."""

    def __init__(self, strategy: str, num_bits: int, group_size: Optional[int] = None):
        if strategy == "group" and group_size is None:
            raise ValueError("group_size must be given when using strategy group")
        self.strategy, self.group_size, self.num_bits, self.tile_size = strategy, group_size, num_bits, 16

    def process_weights_after_loading(self, layer: Module) -> None:
        pass

    def create_weights(self, layer: Module, input_size: int, output_partition_sizes: List[int], input_size_per_partition: int,
                       params_dtype: torch.dtype, weight_loader: Callable, **kwargs):
        pf, n_part, k_part = 32 // self.num_bits, sum(output_partition_sizes), input_size_per_partition
        qweight = Parameter(torch.empty(k_part // self.tile_size // 2, n_part * self.tile_size // pf, dtype=torch.int32), requires_grad=False)
        set_weight_attrs(qweight, {"input_dim": 0, "output_dim": 1, "packed_dim": 1, "pack_factor": pf,
                                   "marlin_tile_size": self.tile_size, "weight_loader": weight_loader})
        layer.register_parameter("weight_packed", qweight)
        groups = 1 if self.group_size is None else k_part // self.group_size
        scales = Parameter(torch.empty(groups, n_part, dtype=params_dtype), requires_grad=False)
        set_weight_attrs(scales, {"output_dim": 1, "input_dim": None if groups == 1 else 0, "weight_loader": weight_loader})
        layer.register_parameter("scale_packed", scales)
        weight_shape = Parameter(torch.empty(2, dtype=torch.int64), requires_grad=False)
        layer.register_parameter("weight_shape", weight_shape)
        set_weight_attrs(weight_shape, {"weight_loader": weight_loader})
        meta = Parameter(torch.empty(k_part // 8 // 2 // 2, n_part * 2, dtype=torch.int16), requires_grad=False)
        set_weight_attrs(meta, {"input_dim": 0, "packed_dim": 1, "pack_factor": 1, "output_dim": 1, "marlin_tile_size": 2,
                                "weight_loader": weight_loader})
        layer.register_parameter("meta", meta)
        layer.workspace = Parameter(torch.zeros((n_part // GPTQ_MARLIN_24_MIN_THREAD_N) * GPTQ_MARLIN_24_MAX_PARALLEL, dtype=torch.int),
                                    requires_grad=False)

    def apply_weights(self, layer: Module, x: torch.Tensor):
        x2 = x.view(-1, x.shape[-1])
        out = ops.gptq_marlin_24_gemm(x2, layer.weight_packed, layer.meta, layer.scale_packed, layer.workspace, self.num_bits,
                                      x2.shape[0], layer.scale_packed.shape[1], x2.shape[1])
        return out.view(x.shape[:-1] + (out.shape[1],))


class CompressedTensorsW8A8(CompressedTensorsScheme):
    """int8 weight [N, K] + weight_scale (per tensor per logical weight, or per channel) (+ static input_scale)."""

    def __init__(self, strategy: str, is_static_input_scheme: bool):
        self.strategy, self.is_static_input_scheme = strategy, is_static_input_scheme

    def process_weights_after_loading(self, layer: Module) -> None:
        # the GEMM takes one scale or one per output channel: a fused layer with per-tensor scales becomes per-channel
        if self.strategy == QuantizationStrategy.TENSOR and len(self.logical_widths) > 1:
            per_channel = torch.empty((sum(self.logical_widths), 1), dtype=torch.float32, device=layer.weight_scale.device)
            start = 0
            for idx, width in enumerate(self.logical_widths):
                per_channel[start:start + width, :] = layer.weight_scale[idx]
                start += width
            layer.weight_scale = Parameter(per_channel, requires_grad=False)
        layer.weight = Parameter(layer.weight.t(), requires_grad=False)   # [K, N] column-major view

    def create_weights(self, layer: Module, output_partition_sizes: List[int], input_size_per_partition: int,
                       params_dtype: torch.dtype, weight_loader: Callable, **kwargs):
        self.logical_widths = output_partition_sizes
        per_channel = self.strategy == QuantizationStrategy.CHANNEL
        shape = (sum(self.logical_widths), 1) if per_channel else (len(self.logical_widths),)
        weight_scale = Parameter(torch.empty(*shape, dtype=torch.float32), requires_grad=False)
        layer.register_parameter("weight_scale", weight_scale)
        set_weight_attrs(weight_scale, {"weight_loader": weight_loader, "output_dim": 0} if per_channel else
                         {"weight_loader": weight_loader, "needs_scalar_to_array": True})
        weight = Parameter(torch.empty(sum(output_partition_sizes), input_size_per_partition, dtype=torch.int8), requires_grad=False)
        layer.register_parameter("weight", weight)
        set_weight_attrs(weight, {"input_dim": 1, "output_dim": 0, "weight_loader": weight_loader})
        if self.is_static_input_scheme:
            input_scale = Parameter(torch.empty(1, dtype=torch.float32), requires_grad=False)
            layer.register_parameter("input_scale", input_scale)
            set_weight_attrs(input_scale, {"weight_loader": weight_loader, "ignore_warning": True})
        else:
            layer.input_scale = None

    def apply_weights(self, layer: Module, x: torch.Tensor):
        x2 = x.reshape(-1, x.shape[-1])
        x_q, x_scale = ops.scaled_int8_quant(x2, layer.input_scale)   # None -> dynamic per-token
        out = ops.cutlass_scaled_mm(x_q, layer.weight, x_scale, layer.weight_scale, x.dtype)
        return out.reshape(x.shape[:-1] + (out.shape[-1],))


class CompressedTensorsConfig(QuantizationConfig):
    def __init__(self, layer_quant_details: Dict[str, Any], ignore: List[str], quant_format: str):
        self.ignore, self.layer_quant_details, self.quant_format = ignore, layer_quant_details, quant_format

    def get_linear_method(self) -> "CompressedTensorsLinearMethod":
        return CompressedTensorsLinearMethod(self)

    @classmethod
    def get_supported_act_dtypes(cls) -> List[torch.dtype]:
        return [torch.float16, torch.bfloat16]

    @classmethod
    def get_min_capability(cls) -> int:
        return 75

    @classmethod
    def get_name(cls) -> str:
        return "compressed_tensors"

    @classmethod
    def get_config_filenames(cls) -> List[str]:
        return []

    def get_quant_method(self, layer: Module) -> Optional["CompressedTensorsLinearMethod"]:
        return CompressedTensorsLinearMethod(self) if isinstance(layer, LinearBase) else None

    @classmethod
    def from_config(cls, config: Dict[str, Any]) -> "CompressedTensorsConfig":
        details: Dict[str, Any] = {}
        for _, group in config["config_groups"].items():
            for target in group.get("targets"):
                entry = {"weights": QuantizationArgs.parse_obj(group.get("weights"))}
                try:
                    entry["input_activations"] = QuantizationArgs.parse_obj(group.get("input_activations"))
                except Exception:
                    entry["input_activations"] = None
                details[target] = entry
        return cls(layer_quant_details=details, ignore=config.get("ignore", None), quant_format=config.get("format", None))

    @staticmethod
    def _w8a8_common(w: QuantizationArgs, a: QuantizationArgs) -> bool:
        return (w.num_bits == a.num_bits == 8 and w.strategy in (QuantizationStrategy.TENSOR.value, QuantizationStrategy.CHANNEL.value)
                and w.symmetric and a.symmetric and not w.dynamic)

    def _is_static_tensor_w8a8(self, w, a) -> bool:
        return self._w8a8_common(w, a) and a.strategy == QuantizationStrategy.TENSOR.value and not a.dynamic

    def _is_dynamic_token_w8a8(self, w, a) -> bool:
        return self._w8a8_common(w, a) and a.strategy == QuantizationStrategy.TOKEN.value and a.dynamic

    def _is_wNa16_group_channel(self, w, a) -> bool:
        return (a is None and w.symmetric and not w.dynamic
                and w.strategy in (QuantizationStrategy.CHANNEL.value, QuantizationStrategy.GROUP.value))

    def _get_schema(self, weight_quant: QuantizationArgs, input_quant: Optional[QuantizationArgs]) -> CompressedTensorsScheme:
        if self._is_wNa16_group_channel(weight_quant, input_quant):
            if self.quant_format == CompressionFormat.marlin_24.value and weight_quant.num_bits in W4A16SPARSE24_SUPPORTED_BITS:
                return CompressedTensorsW4A16Sparse24(strategy=weight_quant.strategy, num_bits=weight_quant.num_bits,
                                                      group_size=weight_quant.group_size)
            if self.quant_format == CompressionFormat.pack_quantized.value and weight_quant.num_bits in WNA16_SUPPORTED_BITS:
                return CompressedTensorsWNA16(num_bits=weight_quant.num_bits, strategy=weight_quant.strategy,
                                              group_size=weight_quant.group_size)
        if self.quant_format == CompressionFormat.int_quantized.value and input_quant is not None:
            if self._is_static_tensor_w8a8(weight_quant, input_quant):
                return CompressedTensorsW8A8(strategy=weight_quant.strategy, is_static_input_scheme=True)
            if self._is_dynamic_token_w8a8(weight_quant, input_quant):
                return CompressedTensorsW8A8(strategy=weight_quant.strategy, is_static_input_scheme=False)
        raise NotImplementedError("No compressed-tensors compatible scheme was found.")

    def get_scheme(self, layer: Module) -> CompressedTensorsScheme:
        name = find_first_name_or_class_match(name="", module=layer, targets=self.layer_quant_details.keys(), check_contains=True)
        if name is None:
            raise ValueError(f"Could not matching target for layer {layer}")
        d = self.layer_quant_details.get(name)
        if d is None:
            raise ValueError(f"Could not find quantization details for {layer}.")
        return self._get_schema(weight_quant=d["weights"], input_quant=d["input_activations"])


class CompressedTensorsLinearMethod(LinearMethodBase):
    def __init__(self, quantization_config: CompressedTensorsConfig):
        self.quantization_config = quantization_config

    def process_weights_after_loading(self, layer: Module) -> None:
        return layer.scheme.process_weights_after_loading(layer)

    def create_weights(self, layer: Module, input_size_per_partition: int, output_partition_sizes: List[int], input_size: int,
                       output_size: int, params_dtype: torch.dtype, **extra_weight_attrs):
        scheme = self.quantization_config.get_scheme(layer=layer)
        scheme.create_weights(layer=layer, input_size=input_size, input_size_per_partition=input_size_per_partition,
                              output_partition_sizes=output_partition_sizes, output_size=output_size, params_dtype=params_dtype,
                              weight_loader=extra_weight_attrs.get("weight_loader"))
        layer.scheme = scheme

    def apply(self, layer: Module, x: torch.Tensor, bias: Optional[torch.Tensor] = None):
        if bias is not None:
            raise ValueError("bias is not supported for this linear method")
        if getattr(layer, "scheme", None) is None:
            raise ValueError("A scheme must be defined for each layer")
        return layer.scheme.apply_weights(layer, x)
