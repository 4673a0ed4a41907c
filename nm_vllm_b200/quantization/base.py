"""Interfaces the quantized linear methods plug into -- the parts of the reference's layer framework the hot path
touches: `set_weight_attrs` (vllm/model_executor/utils.py:15-32), `QuantizeMethodBase` / `QuantizationConfig`
(layers/quantization/base_config.py:8-118) and `LinearMethodBase` (layers/linear.py:31-75).  Only the contract is
mirrored (method names, argument order, the attributes weight loaders read); model loading itself is out of scope."""
from abc import ABC, abstractmethod
from typing import Any, Dict, List, Optional

import torch


def set_weight_attrs(weight: torch.Tensor, weight_attrs: Optional[Dict[str, Any]]) -> None:
    """Attach loader metadata (input_dim / output_dim / packed_dim / pack_factor / marlin_tile_size / weight_loader ...)
    to a parameter; an attribute may be set only once, like the reference."""
    if weight_attrs is None:
        return
    for key, value in weight_attrs.items():
        assert not hasattr(weight, key), f"Overwriting existing tensor attribute: {key}"
        setattr(weight, key, value)


class QuantizeMethodBase(ABC):
    @abstractmethod
    def create_weights(self, layer: torch.nn.Module, *weight_args, **extra_weight_attrs):
        raise NotImplementedError

    @abstractmethod
    def apply(self, layer: torch.nn.Module, *args, **kwargs) -> torch.Tensor:
        raise NotImplementedError

    def process_weights_after_loading(self, layer: torch.nn.Module) -> None:
        return


class LinearMethodBase(QuantizeMethodBase):
    """create_weights(layer, input_size_per_partition, output_partition_sizes, input_size, output_size, params_dtype,
    **extra_weight_attrs); apply(layer, x, bias=None) -- linear.py:31-75."""

    @abstractmethod
    def create_weights(self, layer: torch.nn.Module, input_size_per_partition: int, output_partition_sizes: List[int],
                       input_size: int, output_size: int, params_dtype: torch.dtype, **extra_weight_attrs):
        raise NotImplementedError

    @abstractmethod
    def apply(self, layer: torch.nn.Module, x: torch.Tensor, bias: Optional[torch.Tensor] = None) -> torch.Tensor:
        raise NotImplementedError


class QuantizationConfig(ABC):
    @classmethod
    @abstractmethod
    def get_name(cls) -> str:
        raise NotImplementedError

    @classmethod
    @abstractmethod
    def get_supported_act_dtypes(cls) -> List[torch.dtype]:
        raise NotImplementedError

    @classmethod
    @abstractmethod
    def get_min_capability(cls) -> int:
        raise NotImplementedError

    @classmethod
    @abstractmethod
    def get_config_filenames(cls) -> List[str]:
        raise NotImplementedError

    @classmethod
    @abstractmethod
    def from_config(cls, config: Dict[str, Any]) -> "QuantizationConfig":
        raise NotImplementedError

    @classmethod
    def override_quantization_method(cls, hf_quant_cfg, user_quant) -> Optional[str]:
        return None

    @staticmethod
    def get_from_keys(config: Dict[str, Any], keys: List[str]) -> Any:
        for key in keys:
            if key in config:
                return config[key]
        raise ValueError(f"Cannot find any of {keys} in the model's quantization config.")

    @staticmethod
    def get_from_keys_or(config: Dict[str, Any], keys: List[str], default: Any) -> Any:
        try:
            return QuantizationConfig.get_from_keys(config, keys)
        except ValueError:
            return default

    @abstractmethod
    def get_quant_method(self, layer: torch.nn.Module) -> Optional[QuantizeMethodBase]:
        raise NotImplementedError

    def get_scaled_act_names(self) -> List[str]:
        return []


class LinearBase(torch.nn.Module):
    """Marker base of the layers a LinearMethod attaches to (linear.py:117-145): holds the sizes and the method."""

    def __init__(self, input_size: int, output_size: int, params_dtype: Optional[torch.dtype] = None,
                 quant_config: Optional[QuantizationConfig] = None):
        super().__init__()
        self.input_size, self.output_size = input_size, output_size
        self.params_dtype = params_dtype or torch.get_default_dtype()
        self.quant_method = quant_config.get_quant_method(self) if quant_config is not None else None


def device_capability() -> int:
    if not torch.cuda.is_available():
        return 100   # this build targets sm_100a only
    major, minor = torch.cuda.get_device_capability()
    return major * 10 + minor
