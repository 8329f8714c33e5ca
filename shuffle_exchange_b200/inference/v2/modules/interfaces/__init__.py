"""Abstract module interfaces + their registries (reference ``inference/v2/modules/interfaces/*``)."""
from __future__ import annotations

from abc import abstractmethod
from typing import Any, Dict, Optional, Tuple, Type

import torch

from ...inference_parameter import InferenceParameter
from ...ragged.ragged_wrapper import RaggedBatchWrapper
from ..configs import (DSEmbeddingsConfig, DSLinearConfig, DSMoEConfig, DSNormConfig, DSSelfAttentionConfig,
                       DSUnembedConfig)
from ..ds_module import DSModuleBase
from ..module_registry import DSModuleRegistryBase


def _param(t: torch.Tensor) -> InferenceParameter:
    return t if isinstance(t, InferenceParameter) else InferenceParameter.initialize(t)


# ------------------------------------------------------------------------------------------ attention
class DSSelfAttentionBase(DSModuleBase):
    """Ragged self attention: ``forward(q_k_v [tokens, (hq + 2 hkv) * d], kv_cache, batch) -> [tokens, hq * d]``; the
    implementation applies the positional embedding, appends K/V to the blocked cache and attends causally."""

    @staticmethod
    def config_class() -> Type[DSSelfAttentionConfig]:
        return DSSelfAttentionConfig

    def __init__(self, config: DSSelfAttentionConfig, implementation_config: Dict[str, Any]) -> None:
        super().__init__(config, implementation_config)

    @property
    def kv_block_size(self) -> int:
        raise NotImplementedError

    @property
    def q_block_size(self) -> int:
        raise NotImplementedError

    def build_atoms(self, ragged_batch: RaggedBatchWrapper) -> None:
        """Host-side schedule construction hook (kernels here take the ragged metadata directly)."""

    @abstractmethod
    def forward(self, q_k_v: torch.Tensor, kv_cache: torch.Tensor, batch: RaggedBatchWrapper,
                inv_freqs: Optional[torch.Tensor] = None) -> torch.Tensor:
        ...


class DSSelfAttentionRegistry(DSModuleRegistryBase):
    registry: Dict = {}

    @staticmethod
    def associated_class() -> Type[DSModuleBase]:
        return DSSelfAttentionBase


# ------------------------------------------------------------------------------------------ embedding
class DSEmbeddingBase(DSModuleBase):

    @staticmethod
    def config_class() -> Type[DSEmbeddingsConfig]:
        return DSEmbeddingsConfig

    def transform_param(self, embed_param: torch.Tensor) -> InferenceParameter:
        return _param(embed_param)

    @property
    @abstractmethod
    def output(self) -> torch.Tensor:
        ...

    @abstractmethod
    def forward(self, ragged_batch: RaggedBatchWrapper, word_embeddings: torch.Tensor,
                position_embeddings: Optional[torch.Tensor] = None,
                token_type_ids: Optional[torch.Tensor] = None,
                token_type_embeddings: Optional[torch.Tensor] = None) -> torch.Tensor:
        ...


class DSEmbeddingRegistry(DSModuleRegistryBase):
    registry: Dict = {}

    @staticmethod
    def associated_class() -> Type[DSModuleBase]:
        return DSEmbeddingBase


# ------------------------------------------------------------------------------------------ linear
class DSLinearBase(DSModuleBase):

    @staticmethod
    def config_class() -> Type[DSLinearConfig]:
        return DSLinearConfig

    @abstractmethod
    def transform_param(self, param: torch.Tensor) -> InferenceParameter:
        """Weights ``[out, in]`` (2-D) and biases (1-D) -> the layout / precision the implementation computes in."""

    @abstractmethod
    def forward(self, hidden_states: torch.Tensor, w: torch.Tensor, b: Optional[torch.Tensor] = None) -> torch.Tensor:
        ...

    @property
    @abstractmethod
    def output(self) -> torch.Tensor:
        ...


class DSLinearRegistry(DSModuleRegistryBase):
    registry: Dict = {}

    @staticmethod
    def associated_class() -> Type[DSModuleBase]:
        return DSLinearBase


# ------------------------------------------------------------------------------------------ MoE
class DSMoEBase(DSModuleBase):

    @staticmethod
    def config_class() -> Type[DSMoEConfig]:
        return DSMoEConfig

    @abstractmethod
    def transform_gate_param(self, param: torch.Tensor) -> InferenceParameter:
        ...

    @abstractmethod
    def transform_moe_mlp_1_param(self, param: torch.Tensor) -> InferenceParameter:
        ...

    @abstractmethod
    def transform_moe_mlp_2_param(self, param: torch.Tensor) -> InferenceParameter:
        ...

    @abstractmethod
    def forward(self, hidden_states: torch.Tensor, gate_w: torch.Tensor, mlp_1_w: torch.Tensor,
                mlp_2_w: torch.Tensor, mlp_1_b: Optional[torch.Tensor] = None,
                mlp_2_b: Optional[torch.Tensor] = None) -> torch.Tensor:
        ...

    @property
    @abstractmethod
    def output(self) -> torch.Tensor:
        ...


class DSMoERegistry(DSModuleRegistryBase):
    registry: Dict = {}

    @staticmethod
    def associated_class() -> Type[DSModuleBase]:
        return DSMoEBase


# ------------------------------------------------------------------------------------------ norms
class DSPreNormBase(DSModuleBase):
    """``forward(residual, hidden_states, gamma, beta) -> (residual + hidden_states, norm(residual + hidden_states))``;
    ``hidden_states`` may be ``None`` for the first layer."""

    @staticmethod
    def config_class() -> Type[DSNormConfig]:
        return DSNormConfig

    def transform_param(self, param: torch.Tensor) -> InferenceParameter:
        return _param(param)

    @abstractmethod
    def forward(self, residual: torch.Tensor, hidden_states: Optional[torch.Tensor], gamma: torch.Tensor,
                beta: Optional[torch.Tensor] = None) -> Tuple[torch.Tensor, torch.Tensor]:
        ...


class DSPreNormRegistry(DSModuleRegistryBase):
    registry: Dict = {}

    @staticmethod
    def associated_class() -> Type[DSModuleBase]:
        return DSPreNormBase


class DSPostNormBase(DSModuleBase):
    """``forward(residual, hidden_states, gamma, beta) -> norm(residual + hidden_states)``."""

    @staticmethod
    def config_class() -> Type[DSNormConfig]:
        return DSNormConfig

    def transform_param(self, param: torch.Tensor) -> InferenceParameter:
        return _param(param)

    @abstractmethod
    def forward(self, residual: torch.Tensor, hidden_states: torch.Tensor, gamma: torch.Tensor,
                beta: Optional[torch.Tensor] = None) -> torch.Tensor:
        ...


class DSPostNormRegistry(DSModuleRegistryBase):
    registry: Dict = {}

    @staticmethod
    def associated_class() -> Type[DSModuleBase]:
        return DSPostNormBase


# ------------------------------------------------------------------------------------------ unembed
class DSUnembedBase(DSModuleBase):
    """Final norm (optional) + last-token gather + vocabulary projection -> ``[n_seqs, vocab]`` logits."""

    @staticmethod
    def config_class() -> Type[DSUnembedConfig]:
        return DSUnembedConfig

    def transform_param(self, param: torch.Tensor) -> InferenceParameter:
        return _param(param)

    @property
    @abstractmethod
    def output(self) -> torch.Tensor:
        ...

    @abstractmethod
    def forward(self, hidden_states: torch.Tensor, vocab_embedding: torch.Tensor, ragged_metadata: RaggedBatchWrapper,
                bias: Optional[torch.Tensor] = None, gamma: Optional[torch.Tensor] = None,
                beta: Optional[torch.Tensor] = None) -> torch.Tensor:
        ...


class DSUnembedRegistry(DSModuleRegistryBase):
    registry: Dict = {}

    @staticmethod
    def associated_class() -> Type[DSModuleBase]:
        return DSUnembedBase
