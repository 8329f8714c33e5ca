"""QKV projections (reference: ``common_parameters/qkv_parameters.py``)."""
import torch

from ..parameter_base import ParameterBase


class FusedQKVParameter(ParameterBase):
    """Checkpoint already stores q|k|v rows in order."""
    params: torch.Tensor

    def finalize(self) -> torch.Tensor:
        return self.params


class UnfusedQKVParameter(ParameterBase):
    q_params: torch.Tensor
    k_params: torch.Tensor
    v_params: torch.Tensor

    def finalize(self) -> torch.Tensor:
        return torch.cat([self.q_params, self.k_params, self.v_params], dim=0).contiguous()


def _megatron_regroup(p: torch.Tensor, n_q: int, n_kv: int, d: int) -> torch.Tensor:
    """Rows ordered ``[kv_group, (q_0..q_{r-1}, k, v), d]`` -> q|k|v blocks (``r = n_q / n_kv``)."""
    r = n_q // n_kv
    tail = p.shape[1:]
    g = p.reshape(n_kv, r + 2, d, *tail)
    q = g[:, :r].reshape(n_q * d, *tail)
    k = g[:, r].reshape(n_kv * d, *tail)
    v = g[:, r + 1].reshape(n_kv * d, *tail)
    return torch.cat([q, k, v], dim=0).contiguous()


class MegatronQKVParameter(ParameterBase):
    """MHA, rows ``[heads, (q|k|v), d]`` (BLOOM, GPT-NeoX, Falcon-MHA)."""
    params: torch.Tensor

    def finalize(self) -> torch.Tensor:
        m = self.inference_model
        return _megatron_regroup(self.params, m.n_heads, m.n_heads, m.head_size)


class GQAMegatronQKVParameter(ParameterBase):
    """GQA / MQA, rows grouped per KV head (Falcon new decoder architecture)."""
    params: torch.Tensor

    def finalize(self) -> torch.Tensor:
        m = self.inference_model
        return _megatron_regroup(self.params, m.n_heads, m.n_heads_kv, m.head_size)
