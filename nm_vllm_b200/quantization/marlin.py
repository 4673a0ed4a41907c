"""Pre-packed Marlin checkpoints (dense W4A16 and 2:4-sparse W4A16 / W8A16): mirrors of
vllm/model_executor/layers/quantization/marlin.py (`MarlinConfig` / `MarlinLinearMethod`, fp16, group {-1, 128},
parameters B / s / workspace) and gptq_marlin_24.py (`GPTQMarlin24Config` :25-120, `GPTQMarlin24LinearMethod`
:123-291, parameters B_24 / B_meta / s / workspace).  The tensors arrive in kernel layout from the checkpoint: nothing is
repacked here, the loader attributes (`marlin_tile_size`, `packed_dim`, `pack_factor`) describe how shards are cut."""
from typing import Any, Dict, List, Optional

import torch
from torch.nn.parameter import Parameter

from .. import _custom_ops as ops
from .base import LinearBase, LinearMethodBase, QuantizationConfig, set_weight_attrs

GPTQ_MARLIN_24_TILE = 16
GPTQ_MARLIN_24_MIN_THREAD_N = 128
GPTQ_MARLIN_24_MIN_THREAD_K = 128
GPTQ_MARLIN_24_MAX_PARALLEL = 64
GPTQ_MARLIN_24_SUPPORTED_NUM_BITS = [4, 8]
GPTQ_MARLIN_24_SUPPORTED_GROUP_SIZES = [-1, 128]
GPTQ_MARLIN_24_SUPPORTED_SYM = [True]


def _dev():
    return "cuda" if torch.cuda.is_available() else "cpu"


class _MarlinFamilyConfig(QuantizationConfig):
    @classmethod
    def get_supported_act_dtypes(cls) -> List[torch.dtype]:
        return [torch.half]

    @classmethod
    def get_min_capability(cls) -> int:
        return 80

    @classmethod
    def get_config_filenames(cls) -> List[str]:
        return ["quantize_config.json"]


class MarlinConfig(_MarlinFamilyConfig):
    def __init__(self, group_size: int, lm_head_quantized: bool = False) -> None:
        if group_size not in (128, -1):
            raise ValueError("Currently, only group size 128 and -1 (channelwise) is supported for Marlin, "
                             f"but got group_size of {group_size}")
        self.group_size, self.lm_head_quantized = group_size, lm_head_quantized
        self.pack_factor, self.tile_size = 32 // 4, 16
        self.min_n_threads, self.min_k_threads, self.max_parallel, self.perm_len = 64, 128, 16, 1024

    def __repr__(self) -> str:
        return f"MarlinConfig(group_size={self.group_size}, lm_head_quantized={self.lm_head_quantized})"

    @classmethod
    def get_name(cls) -> str:
        return "marlin"

    @classmethod
    def from_config(cls, config: Dict[str, Any]) -> "MarlinConfig":
        return cls(cls.get_from_keys(config, ["group_size"]), cls.get_from_keys_or(config, ["lm_head"], default=False))

    def get_quant_method(self, layer: torch.nn.Module) -> Optional["MarlinLinearMethod"]:
        if isinstance(layer, LinearBase) or (self.lm_head_quantized and getattr(layer, "is_lm_head", False)):
            return MarlinLinearMethod(self)
        return None


class GPTQMarlin24Config(_MarlinFamilyConfig):
    def __init__(self, weight_bits: int, group_size: int) -> None:
        if weight_bits not in GPTQ_MARLIN_24_SUPPORTED_NUM_BITS:
            raise ValueError(f"Marlin_24 does not support weight_bits = {weight_bits}. "
                             f"Only weight_bits = {GPTQ_MARLIN_24_SUPPORTED_NUM_BITS} are supported.")
        if group_size not in GPTQ_MARLIN_24_SUPPORTED_GROUP_SIZES:
            raise ValueError(f"Marlin_24 does not support group_size = {group_size}. "
                             f"Only group_sizes = {GPTQ_MARLIN_24_SUPPORTED_GROUP_SIZES} are supported.")
        self.weight_bits, self.group_size = weight_bits, group_size
        self.pack_factor, self.tile_size = 32 // weight_bits, GPTQ_MARLIN_24_TILE
        self.min_n_threads, self.min_k_threads = GPTQ_MARLIN_24_MIN_THREAD_N, GPTQ_MARLIN_24_MIN_THREAD_K
        self.max_parallel, self.perm_len = GPTQ_MARLIN_24_MAX_PARALLEL, 1024

    def __repr__(self) -> str:
        return "Marlin24Config(weight_bits={}, group_size={})".format(self.weight_bits, self.group_size)

    @classmethod
    def get_name(cls) -> str:
        return "gptq_marlin_24"

    @classmethod
    def from_config(cls, config: Dict[str, Any]) -> "GPTQMarlin24Config":
        return cls(cls.get_from_keys(config, ["bits"]), cls.get_from_keys(config, ["group_size"]))

    @classmethod
    def override_quantization_method(cls, hf_quant_cfg, user_quant) -> Optional[str]:
        is_24 = hf_quant_cfg.get("checkpoint_format") == "marlin_24"
        if is_24 and (user_quant is None or user_quant in ("gptq", "gptq_marlin_24")):
            return cls.get_name()
        return None

    def get_quant_method(self, layer: torch.nn.Module) -> Optional["GPTQMarlin24LinearMethod"]:
        return GPTQMarlin24LinearMethod(self) if isinstance(layer, LinearBase) else None


def _check_shard(cfg, k_part: int, n_part: int, params_dtype: torch.dtype) -> None:
    if params_dtype != torch.float16:
        raise ValueError(f"The params dtype must be float16, but got {params_dtype}")
    if n_part % cfg.min_n_threads != 0:
        raise ValueError(f"Weight output_size_per_partition = {n_part} is not divisible by min_n_threads = {cfg.min_n_threads}.")
    if n_part % cfg.pack_factor != 0:
        raise ValueError(f"Weight output_size_per_partition = {n_part} is not divisible by pack_factor = {cfg.pack_factor}.")
    if k_part % cfg.min_k_threads != 0:
        raise ValueError(f"Weight input_size_per_partition = {k_part} is not divisible by min_k_threads = {cfg.min_k_threads}.")
    if cfg.group_size != -1 and k_part % cfg.group_size != 0:
        raise ValueError(f"Weight input_size_per_partition = {k_part} is not divisible by group_size = {cfg.group_size}.")
    # a permutation group (1024 values = 4 tiles of 16 x 16) must not straddle two ranks
    if n_part % (cfg.perm_len // cfg.tile_size ** 2) != 0:
        raise ValueError("Each permutation group must reside on the same gpu")


def _scales_and_workspace(cfg, k_part: int, n_part: int, params_dtype):
    groups = 1 if cfg.group_size == -1 else k_part // cfg.group_size
    scales = Parameter(torch.empty(groups, n_part, device=_dev(), dtype=params_dtype), requires_grad=False)
    set_weight_attrs(scales, {"input_dim": None if groups == 1 else 0, "output_dim": 1})
    workspace = Parameter(torch.zeros((n_part // cfg.min_n_threads) * cfg.max_parallel, device=_dev(), dtype=torch.int),
                          requires_grad=False)
    return scales, workspace


class MarlinLinearMethod(LinearMethodBase):
    def __init__(self, quant_config: MarlinConfig):
        self.quant_config = quant_config

    def create_weights(self, layer, input_size_per_partition, output_partition_sizes, input_size, output_size, params_dtype,
                       **extra_weight_attrs):
        del input_size, output_size
        cfg, n_part = self.quant_config, sum(output_partition_sizes)
        _check_shard(cfg, input_size_per_partition, n_part, params_dtype)
        qweight = Parameter(torch.empty(input_size_per_partition // cfg.tile_size, n_part * cfg.tile_size // cfg.pack_factor,
                                        device=_dev(), dtype=torch.int32), requires_grad=False)
        set_weight_attrs(qweight, {"input_dim": 0, "output_dim": 1, "packed_dim": 1, "pack_factor": cfg.pack_factor,
                                   "marlin_tile_size": cfg.tile_size})
        scales, workspace = _scales_and_workspace(cfg, input_size_per_partition, n_part, params_dtype)
        for name, p in (("B", qweight), ("s", scales), ("workspace", workspace)):
            layer.register_parameter(name, p)
            set_weight_attrs(p, extra_weight_attrs)

    def apply(self, layer, x, bias=None):
        x2 = x.view(-1, x.shape[-1])
        out = ops.marlin_gemm(x2, layer.B, layer.s, layer.workspace, x2.shape[0], layer.s.shape[1], x2.shape[1])
        out = out.view(x.shape[:-1] + (out.shape[1],))
        if bias is not None:
            out.add_(bias)
        return out


class GPTQMarlin24LinearMethod(LinearMethodBase):
    def __init__(self, quant_config: GPTQMarlin24Config):
        self.quant_config = quant_config

    def create_weights(self, layer, input_size_per_partition, output_partition_sizes, input_size, output_size, params_dtype,
                       **extra_weight_attrs):
        del input_size, output_size
        cfg, n_part, k_part = self.quant_config, sum(output_partition_sizes), input_size_per_partition
        _check_shard(cfg, k_part, n_part, params_dtype)
        # 2 of every 4 weights along K are kept: K / 2 compressed rows in 16-row tiles
        qweight = Parameter(torch.empty(k_part // cfg.tile_size // 2, n_part * cfg.tile_size // cfg.pack_factor,
                                        device=_dev(), dtype=torch.int32), requires_grad=False)
        set_weight_attrs(qweight, {"input_dim": 0, "output_dim": 1, "packed_dim": 1, "pack_factor": cfg.pack_factor,
                                   "marlin_tile_size": cfg.tile_size})
        # 2-bit position indices: 8 per int16, as [K / 32, 2 N]
        meta = Parameter(torch.empty(k_part // 8 // 2 // 2, n_part * 2, device=_dev(), dtype=torch.int16), requires_grad=False)
        set_weight_attrs(meta, {"input_dim": 0, "packed_dim": 1, "pack_factor": 1, "output_dim": 1, "marlin_tile_size": 2})
        scales, workspace = _scales_and_workspace(cfg, k_part, n_part, params_dtype)
        for name, p in (("B_24", qweight), ("B_meta", meta), ("s", scales), ("workspace", workspace)):
            layer.register_parameter(name, p)
            set_weight_attrs(p, extra_weight_attrs)

    def apply(self, layer, x, bias=None):
        x2 = x.view(-1, x.shape[-1])
        out = ops.gptq_marlin_24_gemm(x2, layer.B_24, layer.B_meta, layer.s, layer.workspace, self.quant_config.weight_bits,
                                      x2.shape[0], layer.s.shape[1], x2.shape[1])
        out = out.view(x.shape[:-1] + (out.shape[1],))
        if bias is not None:
            out.add_(bias)
        return out
