"""Fill an (already injected or AutoTP-sharded) model straight from checkpoint state dicts, layer by layer, without
materialising the full model first (reference: ``module_inject/load_checkpoint.py:1-270``).

The reference walks the module tree with a per-class policy table; here the walk is generic: every parameter / buffer is
looked up under its qualified name (plus ``prefix`` candidates), tensor-parallel slicing is delegated to the target
module when it knows how (``AutoTP`` layers expose ``shard_from_full``), otherwise to ``mp_replace.copy`` which slices
along whichever dimension differs.  Injected fused layers (:class:`HFLayerAdapter`) load through their container policy:
the client layer is materialised from the state dict on the meta device's shapes, then re-injected."""
from __future__ import annotations

import gc
from typing import Dict, Iterable, List, Optional

import torch
import torch.nn as nn

from ..utils.logging import logger


def _candidates(name: str, prefixes: Iterable[str]) -> List[str]:
    out = [name]
    for p in prefixes:
        out.append(p + name)
        if name.startswith(p):
            out.append(name[len(p):])
    return out


def _fit(dst: torch.Tensor, src: torch.Tensor, mp_replace, name: str) -> Optional[torch.Tensor]:
    if tuple(dst.shape) == tuple(src.shape):
        return src
    if mp_replace is not None and dst.dim() == src.dim():
        try:
            return mp_replace.copy(torch.empty_like(dst), src)
        except Exception as e:          # pragma: no cover - shape that no slicing explains
            logger.warning(f"load_model_with_checkpoint: cannot slice {name} {tuple(src.shape)} -> {tuple(dst.shape)}: {e}")
    return None


def load_model_with_checkpoint(r_module: nn.Module, sd, mp_replace=None, ckpt_type: str = "pp", ckpt_mp_size: int = 1,
                               weight_quantizer=None, rank: int = 0, container=None, prefixes=("", "model.",
                                                                                                 "transformer.",
                                                                                                 "module.")):
    """``sd``: one state dict or a list of them (``ckpt_type == "tp"``: shards of one tensor-parallel checkpoint that are
    concatenated on the fly; ``"pp"``: disjoint pieces).  Returns the list of parameter names that were NOT found."""
    sds: List[Dict[str, torch.Tensor]] = list(sd) if isinstance(sd, (list, tuple)) else [sd]
    missing: List[str] = []

    def lookup(name: str, dst: torch.Tensor) -> Optional[torch.Tensor]:
        for cand in _candidates(name, prefixes):
            hits = [s[cand] for s in sds if cand in s]
            if not hits:
                continue
            if len(hits) == 1 or ckpt_type != "tp" or tuple(hits[0].shape) == tuple(dst.shape):
                return hits[0]                       # single piece, or replicated across the shards
            # tensor-parallel shards: concatenate along the first dimension that is too small in a single shard
            dim = next((d for d in range(hits[0].dim()) if hits[0].shape[d] * len(hits) >= dst.shape[d]
                        and hits[0].shape[d] != dst.shape[d]), 0)
            return torch.cat(hits, dim=dim)
        return None

    with torch.no_grad():
        for name, p in list(r_module.named_parameters()) + list(r_module.named_buffers()):
            src = lookup(name, p)
            if src is None:
                missing.append(name)
                continue
            if weight_quantizer is not None and src.dim() == 2 and p.dtype == torch.int8:
                src = weight_quantizer.quantize(src)
            fitted = _fit(p, src, mp_replace, name)
            if fitted is None:
                missing.append(name)
                continue
            if p.device.type == "meta":
                # materialise: swap the meta parameter for a real one on the current device
                dev = torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu")
                owner, _, leaf = name.rpartition(".")
                mod = r_module.get_submodule(owner) if owner else r_module
                new = fitted.to(device=dev, dtype=p.dtype).clone()
                if isinstance(p, nn.Parameter):
                    mod._parameters[leaf] = nn.Parameter(new, requires_grad=p.requires_grad)
                else:
                    mod._buffers[leaf] = new
            else:
                p.data.copy_(fitted.to(p.dtype))
    # tied output embeddings (reference ``load_checkpoint.py:252-268``)
    emb = next((m for n, m in r_module.named_modules() if isinstance(m, nn.Embedding)), None)
    head = getattr(r_module, "lm_head", None) or getattr(r_module, "embed_out", None)
    if emb is not None and isinstance(head, nn.Linear) and any(n.startswith(("lm_head", "embed_out")) for n in missing) \
            and head.weight.shape == emb.weight.shape:
        head.weight = emb.weight
        missing = [n for n in missing if not n.startswith(("lm_head", "embed_out"))]
    for s in sds:
        del s
    gc.collect()
    return missing
