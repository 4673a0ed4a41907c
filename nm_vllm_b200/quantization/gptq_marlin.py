"""GPTQ checkpoints served by the Marlin-layout W4A16 / W8A16 GEMM: mirror of
vllm/model_executor/layers/quantization/gptq_marlin.py (`GPTQMarlinConfig` :59-187, `GPTQMarlinState` :190-192,
`GPTQMarlinLinearMethod` :195-466).  Same parameter names / shapes / loader attributes, same lazy REPACK -> READY step on
the first `apply` (GPTQ qweight -> Marlin tiles through our own `gptq_marlin_repack`, scales through
`marlin_permute_scales`, act-order rows sorted by group), same operator calls."""
import enum
from typing import Any, Dict, List, Optional

import torch
from torch.nn.parameter import Parameter

from .. import _custom_ops as ops
from .base import LinearBase, LinearMethodBase, QuantizationConfig, device_capability, set_weight_attrs
from .layouts import (GPTQ_MARLIN_MAX_PARALLEL, GPTQ_MARLIN_MIN_THREAD_K, GPTQ_MARLIN_MIN_THREAD_N,
                      GPTQ_MARLIN_SUPPORTED_GROUP_SIZES, GPTQ_MARLIN_SUPPORTED_NUM_BITS, GPTQ_MARLIN_SUPPORTED_SYM,
                      GPTQ_MARLIN_TILE, get_pack_factor, marlin_permute_scales)


class GPTQMarlinConfig(QuantizationConfig):
    def __init__(self, weight_bits: int, group_size: int, desc_act: bool, is_sym: bool, lm_head_quantized: bool) -> None:
        if desc_act and group_size == -1:
            desc_act = False   # one group per output channel: act-order changes nothing (gptq_marlin.py:64-67)
        self.weight_bits, self.group_size, self.desc_act = weight_bits, group_size, desc_act
        self.is_sym, self.lm_head_quantized = is_sym, lm_head_quantized
        if weight_bits not in GPTQ_MARLIN_SUPPORTED_NUM_BITS:
            raise ValueError(f"Marlin does not support weight_bits = {weight_bits}. "
                             f"Only weight_bits = {GPTQ_MARLIN_SUPPORTED_NUM_BITS} are supported.")
        if group_size not in GPTQ_MARLIN_SUPPORTED_GROUP_SIZES:
            raise ValueError(f"Marlin does not support group_size = {group_size}. "
                             f"Only group_sizes = {GPTQ_MARLIN_SUPPORTED_GROUP_SIZES} are supported.")
        if is_sym not in GPTQ_MARLIN_SUPPORTED_SYM:
            raise ValueError(f"Marlin does not support is_sym = {is_sym}. Only sym = {GPTQ_MARLIN_SUPPORTED_SYM} are supported.")
        self.pack_factor = get_pack_factor(weight_bits)
        self.tile_size = GPTQ_MARLIN_TILE
        self.min_thread_n = GPTQ_MARLIN_MIN_THREAD_N
        self.min_thread_k = GPTQ_MARLIN_MIN_THREAD_K
        self.max_parallel = GPTQ_MARLIN_MAX_PARALLEL

    def __repr__(self) -> str:
        return (f"GPTQMarlinConfig(weight_bits={self.weight_bits}, group_size={self.group_size}, "
                f"desc_act={self.desc_act}, lm_head_quantized={self.lm_head_quantized})")

    @classmethod
    def get_name(cls) -> str:
        return "gptq_marlin"

    @classmethod
    def get_supported_act_dtypes(cls) -> List[torch.dtype]:
        return [torch.half, torch.bfloat16]

    @classmethod
    def get_min_capability(cls) -> int:
        return 80

    @classmethod
    def get_config_filenames(cls) -> List[str]:
        return ["quantize_config.json"]

    @classmethod
    def from_config(cls, config: Dict[str, Any]) -> "GPTQMarlinConfig":
        return cls(cls.get_from_keys(config, ["bits"]), cls.get_from_keys(config, ["group_size"]),
                   cls.get_from_keys(config, ["desc_act"]), cls.get_from_keys(config, ["sym"]),
                   cls.get_from_keys_or(config, ["lm_head"], default=False))

    @classmethod
    def override_quantization_method(cls, hf_quant_cfg, user_quant) -> Optional[str]:
        if cls.is_marlin_compatible(hf_quant_cfg) and (user_quant is None or user_quant == "marlin"):
            return cls.get_name()
        return None

    def get_quant_method(self, layer: torch.nn.Module) -> Optional["GPTQMarlinLinearMethod"]:
        if isinstance(layer, LinearBase) or (self.lm_head_quantized and getattr(layer, "is_lm_head", False)):
            return GPTQMarlinLinearMethod(self)
        return None

    @classmethod
    def is_marlin_compatible(cls, quant_config: Dict[str, Any]) -> bool:
        bits, group, sym, desc = (quant_config.get(k) for k in ("bits", "group_size", "sym", "desc_act"))
        if bits is None or group is None or sym is None or desc is None:
            return False
        if device_capability() < cls.get_min_capability():
            return False
        return (bits in GPTQ_MARLIN_SUPPORTED_NUM_BITS and group in GPTQ_MARLIN_SUPPORTED_GROUP_SIZES
                and sym in GPTQ_MARLIN_SUPPORTED_SYM)


class GPTQMarlinState(enum.Enum):
    REPACK = enum.auto()
    READY = enum.auto()


def replace_tensor(layer: torch.nn.Module, name: str, new_t: torch.Tensor) -> None:
    """Put a freshly produced tensor into the storage of an already registered parameter (resize_ + copy_ keep the
    registration and reuse the buffer, gptq_marlin.py:389-396)."""
    getattr(layer, name).resize_(new_t.shape)
    getattr(layer, name).copy_(new_t)


class GPTQMarlinLinearMethod(LinearMethodBase):
    def __init__(self, quant_config: GPTQMarlinConfig) -> None:
        self.quant_config = quant_config

    def create_weights(self, layer: torch.nn.Module, input_size_per_partition: int, output_partition_sizes: List[int],
                       input_size: int, output_size: int, params_dtype: torch.dtype, **extra_weight_attrs) -> None:
        del output_size
        cfg = self.quant_config
        group_size = cfg.group_size if cfg.group_size != -1 else input_size
        if params_dtype not in (torch.float16, torch.bfloat16):
            raise ValueError(f"The params dtype must be float16 or bfloat16, but got {params_dtype}")
        n_part = sum(output_partition_sizes)
        if n_part % cfg.min_thread_n != 0:
            raise ValueError(f"Weight output_size_per_partition = {n_part} is not divisible by  min_thread_n = {cfg.min_thread_n}.")
        if input_size_per_partition % cfg.min_thread_k != 0:
            raise ValueError(f"Weight input_size_per_partition = {input_size_per_partition} is not divisible "
                             f"by min_thread_k = {cfg.min_thread_k}.")
        if group_size < input_size and input_size_per_partition % group_size != 0:
            raise ValueError(f"Weight input_size_per_partition = {input_size_per_partition}"
                             f" is not divisible by group_size = {group_size}.")
        # which dimension the scales shard along under tensor parallelism (gptq_marlin.py:246-269): with act-order every
        # rank keeps ALL groups (rows of any group can live in any K shard) and the kernel is told whether it sees all of K
        n_scale_rows, scale_input_dim = input_size // group_size, None
        if cfg.desc_act:
            assert cfg.group_size != -1
            is_k_full = input_size_per_partition == input_size
        else:
            is_k_full = True
            if input_size != input_size_per_partition and cfg.group_size != -1:
                n_scale_rows, scale_input_dim = input_size_per_partition // group_size, 0

        qweight = Parameter(torch.empty(input_size_per_partition // cfg.pack_factor, n_part, dtype=torch.int32), requires_grad=False)
        set_weight_attrs(qweight, {**extra_weight_attrs, "input_dim": 0, "output_dim": 1, "packed_dim": 0,
                                   "pack_factor": cfg.pack_factor})
        g_idx = Parameter(torch.empty(input_size_per_partition, dtype=torch.int32), requires_grad=False)
        set_weight_attrs(g_idx, {**extra_weight_attrs, "input_dim": 0, "ignore_warning": True})
        scales = Parameter(torch.empty(n_scale_rows, n_part, dtype=params_dtype), requires_grad=False)
        set_weight_attrs(scales, {**extra_weight_attrs, "input_dim": scale_input_dim, "output_dim": 1})
        # symmetric only: the zero points are never read, the parameter exists so that checkpoints load (meta device)
        qzeros = Parameter(torch.empty(n_scale_rows, n_part // cfg.pack_factor, dtype=torch.int32, device="meta"), requires_grad=False)
        set_weight_attrs(qzeros, {**extra_weight_attrs, "input_dim": scale_input_dim, "output_dim": 1, "packed_dim": 1,
                                  "pack_factor": cfg.pack_factor})
        layer.register_parameter("qweight", qweight)
        layer.register_parameter("g_idx", g_idx)
        layer.register_parameter("scales", scales)
        layer.register_parameter("qzeros", qzeros)
        layer.g_idx_sort_indices = torch.empty(g_idx.shape, dtype=torch.int32)
        layer.workspace = torch.zeros((n_part // cfg.min_thread_n) * cfg.max_parallel, dtype=torch.int, requires_grad=False)
        layer.input_size_per_partition = input_size_per_partition
        layer.output_size_per_partition = n_part
        layer.input_size = input_size
        layer.is_k_full = is_k_full
        layer.marlin_state = GPTQMarlinState.REPACK

    def _repack(self, layer: torch.nn.Module) -> None:
        cfg = self.quant_config
        dev = layer.qweight.device
        k_part, n_part = layer.input_size_per_partition, layer.output_size_per_partition
        if cfg.desc_act:
            order = torch.argsort(layer.g_idx).to(torch.int)
            replace_tensor(layer, "g_idx", layer.g_idx[order])
            layer.g_idx_sort_indices = layer.g_idx_sort_indices.to(dev)
            replace_tensor(layer, "g_idx_sort_indices", order)
        else:
            layer.g_idx = Parameter(torch.empty(0, dtype=torch.int, device=dev), requires_grad=False)
            layer.g_idx_sort_indices = Parameter(torch.empty(0, dtype=torch.int, device=dev), requires_grad=False)
        replace_tensor(layer, "qweight", ops.gptq_marlin_repack(layer.qweight, layer.g_idx_sort_indices, k_part, n_part, cfg.weight_bits))
        scales_k = layer.input_size if cfg.desc_act else k_part
        replace_tensor(layer, "scales", marlin_permute_scales(layer.scales, scales_k, n_part, cfg.group_size, cfg.weight_bits))
        layer.workspace = layer.workspace.to(dev)

    def apply(self, layer: torch.nn.Module, x: torch.Tensor, bias: Optional[torch.Tensor] = None) -> torch.Tensor:
        x2 = x.reshape(-1, x.shape[-1])
        n_part, k_part = layer.output_size_per_partition, layer.input_size_per_partition
        if layer.marlin_state == GPTQMarlinState.REPACK:
            layer.marlin_state = GPTQMarlinState.READY
            self._repack(layer)
        out = ops.gptq_marlin_gemm(x2, layer.qweight, layer.scales, layer.g_idx, layer.g_idx_sort_indices, layer.workspace,
                                   self.quant_config.weight_bits, x2.shape[0], n_part, k_part, layer.is_k_full)
        if bias is not None:
            out.add_(bias)
        return out.reshape(x.shape[:-1] + (n_part,))

    def can_fuse_silu_and_mul(self, layer: torch.nn.Module) -> bool:
        """gate_up_proj layers whose SiluAndMul can run in the GEMM epilogue (one launch instead of two)."""
        n_part = layer.output_size_per_partition
        if self.quant_config.desc_act or n_part % 128 != 0:
            return False
        # The fused epilogue lives on the stream-K path of the kernel.  A layer with few tiles (a TP shard: tiles * 2 <= SMs)
        # takes the cluster split-K path instead, which is 1.5-2x faster than stream-K there (profiles/r2_gemm_cluster_vs_streamk.txt)
        # -- measured: fusing on such shards made the TP4 / TP8 step slower (2.31 -> 2.43 ms), so they keep the two launches.
        sms = torch.cuda.get_device_properties(layer.qweight.device).multi_processor_count
        return (n_part // 128) * 2 > sms

    def apply_silu_and_mul(self, layer: torch.nn.Module, x: torch.Tensor) -> torch.Tensor:
        """`SiluAndMul()(self.apply(layer, x))` for a merged gate_up layer (llama.py:74-90) in one kernel launch:
        returns [..., n_part // 2], bit-identical to the two-op chain, same stored tensors."""
        x2 = x.reshape(-1, x.shape[-1])
        n_part, k_part = layer.output_size_per_partition, layer.input_size_per_partition
        if layer.marlin_state == GPTQMarlinState.REPACK:
            layer.marlin_state = GPTQMarlinState.READY
            self._repack(layer)
        out = ops.gptq_marlin_gemm_silu_and_mul(x2, layer.qweight, layer.scales, layer.workspace,
                                                self.quant_config.weight_bits, x2.shape[0], n_part, k_part)
        return out.reshape(x.shape[:-1] + (n_part // 2,))
