"""One contiguous byte buffer + a JSON-able table of contents for a built v2 model's (local) weights.

Reference: ``inference/v2/model_implementations/flat_model_helpers.py:19-282`` (``flatten_inference_model`` walks the
layer containers, copies every ``InferenceParameter`` and its auxiliary tensors into a single allocation at
256-byte-aligned offsets and records ``ModelMetadata``; ``restore_inference_model`` re-creates the views).  The same
contract is kept — a single allocation (one ``cudaMalloc``, one file write, one H2D copy on reload) and parameters that
are views into it — over this framework's weight dictionary instead of container objects: entries are addressed by
path (``"embed"``, ``"layers.3.wqkv"``) and auxiliary tensors by ``path#aux``.
"""
from __future__ import annotations

import json
import os
from dataclasses import asdict, dataclass, field
from typing import Dict, Iterator, List, Optional, Tuple

import torch

from ..inference_parameter import STR_TO_DTYPE, InferenceParameter

ALIGNMENT = 256


def pad_to_aligned_offset(offset: int, alignment: int = ALIGNMENT) -> int:
    return (offset + alignment - 1) // alignment * alignment


@dataclass
class TensorMetadata:
    dtype: str
    shape: List[int]
    strides: List[int]
    offset: int


@dataclass
class ParameterMetadata:
    core_param: Optional[TensorMetadata] = None
    aux_params: Dict[str, TensorMetadata] = field(default_factory=dict)


@dataclass
class ModelMetadata:
    params: Dict[str, ParameterMetadata] = field(default_factory=dict)
    n_layers: int = 0
    total_bytes: int = 0
    tp_rank: int = 0
    tp_size: int = 1

    def json(self) -> str:
        return json.dumps(asdict(self))

    @staticmethod
    def parse_raw(text: str) -> "ModelMetadata":
        raw = json.loads(text)
        md = ModelMetadata(n_layers=raw["n_layers"], total_bytes=raw["total_bytes"], tp_rank=raw.get("tp_rank", 0),
                           tp_size=raw.get("tp_size", 1))
        for name, pm in raw["params"].items():
            md.params[name] = ParameterMetadata(
                core_param=TensorMetadata(**pm["core_param"]) if pm["core_param"] else None,
                aux_params={k: TensorMetadata(**v) for k, v in pm["aux_params"].items()})
        return md


def make_param_filename(base: str, rank: int, n_ranks: int) -> str:
    return os.path.join(base, f"params_rank_{rank}_of_{n_ranks}.pt")


def make_metadata_filename(base: str, rank: int, n_ranks: int) -> str:
    return os.path.join(base, f"metadata_rank_{rank}_of_{n_ranks}.json")


def make_model_config_filename(base: str) -> str:
    return os.path.join(base, "ds_model_config.pkl")


def _entries(weights: Dict) -> Iterator[Tuple[str, torch.Tensor]]:
    for k, v in weights.items():
        if k.startswith("_") or k == "layers" or not isinstance(v, torch.Tensor):
            continue
        yield k, v
    for li, lw in enumerate(weights.get("layers", ())):
        for k, v in lw.items():
            if isinstance(v, torch.Tensor):
                yield f"layers.{li}.{k}", v


def _tensor_md(t: torch.Tensor, offset: int) -> TensorMetadata:
    return TensorMetadata(dtype=str(t.dtype), shape=list(t.shape), strides=list(t.stride()), offset=offset)


def flatten_inference_model(weights: Dict, tp_rank: int = 0, tp_size: int = 1,
                            device=None) -> Tuple[torch.Tensor, ModelMetadata]:
    """Returns ``(buffer, metadata)`` and REBINDS every tensor in ``weights`` to a view of ``buffer`` (so the separate
    allocations can be freed).  Tied tensors (same storage and offset, e.g. embed/unembed) are stored once."""
    md = ModelMetadata(n_layers=len(weights.get("layers", ())), tp_rank=tp_rank, tp_size=tp_size)
    plan: List[Tuple[str, Optional[str], torch.Tensor, int]] = []
    seen: Dict[Tuple[int, int, Tuple[int, ...], torch.dtype], int] = {}
    off = 0

    def place(t: torch.Tensor) -> int:
        nonlocal off
        key = (t.untyped_storage().data_ptr(), t.storage_offset(), tuple(t.shape), t.dtype)
        if t.numel() and key in seen:
            return seen[key]
        off = pad_to_aligned_offset(off)
        seen[key] = at = off
        off += t.numel() * t.element_size()
        return at

    for name, t in _entries(weights):
        core = t.as_subclass(torch.Tensor).contiguous()
        at = place(core)
        pm = ParameterMetadata(core_param=_tensor_md(core, at))
        plan.append((name, None, core, at))
        for aux_name, aux in getattr(t, "aux_attrs", {}).items():
            aux = aux.contiguous()
            a_at = place(aux)
            pm.aux_params[aux_name] = _tensor_md(aux, a_at)
            plan.append((name, aux_name, aux, a_at))
        md.params[name] = pm
    md.total_bytes = pad_to_aligned_offset(off)
    first = next((t for _, _, t, _ in plan), None)
    dev = device if device is not None else (first.device if first is not None else "cpu")
    buf = torch.empty(md.total_bytes, dtype=torch.uint8, device=dev)
    for _, _, t, at in plan:
        n = t.numel() * t.element_size()
        if n:
            buf[at:at + n].copy_(t.reshape(-1).view(torch.uint8).to(dev))
    restore_inference_model(buf, md, weights)
    return buf, md


def _view(buf: torch.Tensor, tm: TensorMetadata) -> torch.Tensor:
    dt = STR_TO_DTYPE[tm.dtype]
    n = 1
    for s in tm.shape:
        n *= s
    nbytes = n * torch.tensor([], dtype=dt).element_size()
    return buf[tm.offset:tm.offset + nbytes].view(dt).view(tm.shape)


def restore_inference_model(buffer: torch.Tensor, metadata: ModelMetadata, weights: Optional[Dict] = None) -> Dict:
    """Populate (or create) a weight dictionary whose tensors are views into ``buffer``."""
    from .inference_transformer_base import LayerWeights
    if weights is None:
        weights = {}
    layers = weights.setdefault("layers", [])
    while len(layers) < metadata.n_layers:
        layers.append(LayerWeights())
    for name, pm in metadata.params.items():
        core = _view(buffer, pm.core_param)
        if pm.aux_params:
            t = InferenceParameter.initialize(core, **{k: _view(buffer, v) for k, v in pm.aux_params.items()})
        else:
            t = InferenceParameter(core)
        if name.startswith("layers."):
            _, li, key = name.split(".", 2)
            layers[int(li)][key] = t
        else:
            weights[name] = t
    return weights
