"""``InferenceEngine`` (v1): wrap a PyTorch model for tensor-parallel, low-precision inference.

Reference: ``deepspeed/inference/engine.py:40-625``.  Kept: config handling, TP group creation (or adoption of an
``mpu``), dtype conversion, AutoTP sharding (``replace_with_kernel_inject`` selects the fused path, which in this
framework is the ragged v2 engine — see ``inference/v2``), sharded checkpoint loading, optional CUDA-graph capture of
the forward, ``forward`` / ``generate`` delegation, model-time profiling.
"""
from __future__ import annotations

import os
import time
from typing import Optional

import torch
import torch.nn as nn

from .. import comm as dist
from ..utils.logging import log_dist, logger
from .config import DeepSpeedInferenceConfig

INFERENCE_MODEL_TIMER = "model-forward-inference"


class InferenceEngine(nn.Module):
    inference_mp_group = None
    inference_ep_group = None
    expert_mp_group = None

    def __init__(self, model, config: DeepSpeedInferenceConfig):
        super().__init__()
        self.module = model
        self._config = config
        self._get_model_config_generate(config)
        if hasattr(self.module, "generate"):
            self.generate = self._generate
        if hasattr(self.module, "config"):
            self.config = self.module.config
        self.injection_dict = config.injection_policy
        self.mp_group = config.tensor_parallel.tp_group
        self.mpu = config.tensor_parallel.mpu
        self.quantize_merge_count = 1
        self.quantization_scales = None
        self.ep_group = None
        self.expert_mp_group = None
        self.cuda_graph_created = False
        self.checkpoint_engine = None
        self._model_times = []
        self.model_profile_enabled = False
        self.use_cuda_events = torch.cuda.is_available()
        self._cuda_graphs = None
        self.local_cuda_graph = False
        if config.enable_cuda_graph and not torch.cuda.is_available():
            raise ValueError("enable_cuda_graph needs a CUDA device")
        if config.tensor_parallel.tp_size > 1:
            self._create_model_parallel_group(config)
            config.tensor_parallel.tp_group = self.mp_group
        if isinstance(self.module, nn.Module):
            moe, _ = self._has_moe()
            if moe and dist.is_initialized() and dist.get_world_size() > 1:
                self._create_ep_parallel_group(config.moe.moe_experts if hasattr(config.moe, "moe_experts") else [1])
        self._convert_dtype(config)
        if config.tensor_parallel.tp_size > 1 or config.replace_with_kernel_inject:
            self._apply_injection_policy(config)
        if config.checkpoint is not None:
            self._load_checkpoint(config.checkpoint)
        if config.save_mp_checkpoint_path is not None:
            self._save_mp_checkpoint(config.save_mp_checkpoint_path)
        post_q = getattr(getattr(config.quant, "weight", None), "post_init_quant", None)
        if post_q:
            # group-wise INT4/INT8 weights, converted layer by layer before the move to the device
            from .quantization import _init_group_wise_weight_quantization
            self.module = _init_group_wise_weight_quantization(self.module,
                                                               {"weight_quantization": {"post_init_quant": post_q}})
        dev = torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu")
        if not config.keep_module_on_host:
            self.module.to(dev)
        if config.tensor_parallel.tp_size > 1:
            # identical sampling across the TP group
            seed = torch.tensor([torch.initial_seed() % (2 ** 31)], dtype=torch.int64, device=dev)
            dist.broadcast(seed, src=dist.get_global_rank(self.mp_group, 0) if self.mp_group is not None else 0,
                           group=self.mp_group)
            torch.manual_seed(int(seed.item()))
        self.module.eval()

    # ------------------------------------------------------------------------------ profiling
    def profile_model_time(self, use_cuda_events=True):
        if not self.model_profile_enabled and not self._config.enable_cuda_graph:
            self.module.register_forward_pre_hook(self._pre_forward_hook)
            self.module.register_forward_hook(self._post_forward_hook)
        self.model_profile_enabled = True
        self.use_cuda_events = use_cuda_events and torch.cuda.is_available()
        if self.use_cuda_events:
            self._t0, self._t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    def _pre_forward_hook(self, module, *inputs, **kwargs):
        if self.use_cuda_events:
            self._t0.record()
        else:
            self._start = time.time()

    def _post_forward_hook(self, module, input, output):
        if self.use_cuda_events:
            self._t1.record()
            torch.cuda.synchronize()
            self._model_times.append(self._t0.elapsed_time(self._t1))
        else:
            self._model_times.append((time.time() - self._start) * 1e3)

    def model_times(self):
        assert self.model_profile_enabled, "model profiling is not enabled"
        times, self._model_times = self._model_times, []
        if self._config.enable_cuda_graph and not times:
            raise ValueError("Model times are empty and cuda graph is enabled. If this is a GPT-style model this combo is not supported.")
        return times

    # ---------------------------------------------------------------------------------- groups
    def _get_model_config_generate(self, config):
        self.config = getattr(self.module, "config", None) if config.config is None else config.config

    def _create_model_parallel_group(self, config):
        if InferenceEngine.inference_mp_group is None:
            if not dist.is_initialized():
                dist.init_distributed()
            local_rank = int(os.getenv("LOCAL_RANK", "0"))
            if torch.cuda.is_available():
                torch.cuda.set_device(local_rank)
            if config.tensor_parallel.mpu is not None:
                mpu = config.tensor_parallel.mpu
                self.mp_group = mpu.get_model_parallel_group() if hasattr(mpu, "get_model_parallel_group") else mpu.get_tensor_model_parallel_group()
            else:
                tp, world, rank = config.tensor_parallel.tp_size, dist.get_world_size(), dist.get_rank()
                assert world % tp == 0, f"world size {world} is not divisible by tp_size {tp}"
                for start in range(0, world, tp):
                    g = dist.new_group(list(range(start, start + tp)))
                    if start <= rank < start + tp:
                        self.mp_group = g
            InferenceEngine.inference_mp_group = self.mp_group
        else:
            self.mp_group = InferenceEngine.inference_mp_group

    def _create_ep_parallel_group(self, moe_experts):
        self.ep_group, self.expert_mp_group = {}, {}
        world, rank = dist.get_world_size(), dist.get_rank()
        moe_experts = moe_experts if isinstance(moe_experts, list) else [moe_experts]
        for e in moe_experts:
            size = min(e, world)
            for start in range(0, world, size):
                g = dist.new_group(list(range(start, start + size)))
                if start <= rank < start + size:
                    self.ep_group[e] = g
            self.expert_mp_group[e] = None

    def _has_moe(self):
        from ..moe.layer import MoE
        n = [m.num_experts for m in self.module.modules() if isinstance(m, MoE)]
        return bool(n), n

    # ------------------------------------------------------------------------ dtype / injection
    def _convert_dtype(self, config):
        if not isinstance(self.module, nn.Module):
            return
        if config.dtype == torch.float16:
            self.module.half()
        elif config.dtype == torch.bfloat16:
            self.module.bfloat16()
        elif config.dtype == torch.float32:
            self.module.float()

    def _apply_injection_policy(self, config, client_module=None):
        from ..module_inject import replace_transformer_layer
        checkpoint = None
        if isinstance(self.module, nn.Module):
            if config.injection_policy:
                for cls, names in config.injection_policy.items():
                    config.injection_policy_tuple = tuple(names) if isinstance(names, (list, tuple)) else (names,)
                    replace_transformer_layer(cls, self.module, checkpoint, config, self.config)
            else:
                replace_transformer_layer(None, self.module, checkpoint, config, self.config)

    # ------------------------------------------------------------------------------ checkpoints
    def _load_checkpoint(self, load_dir, load_module_strict=True, tag=None):
        """``load_dir``: a folder with ``*.pt`` / ``*.bin`` / safetensors shards, or a JSON descriptor
        ``{"type": ..., "checkpoints": [...], "version": ...}`` (reference :433-506)."""
        import glob
        import json
        files = []
        desc = load_dir if isinstance(load_dir, dict) else None
        if desc is None and os.path.isfile(load_dir) and load_dir.endswith(".json"):
            with open(load_dir) as f:
                desc = json.load(f)
            desc.setdefault("base_dir", os.path.dirname(load_dir))
        if desc is not None and isinstance(desc.get("checkpoints"), dict):
            return self._load_mp_checkpoint(desc)
        if isinstance(load_dir, dict):
            files = load_dir.get("checkpoints", [])
            base = load_dir.get("base_dir", self._config.base_dir)
            files = [os.path.join(base, f) for f in files]
        elif os.path.isfile(load_dir) and load_dir.endswith(".json"):
            with open(load_dir) as f:
                d = json.load(f)
            base = d.get("base_dir", os.path.dirname(load_dir))
            files = [os.path.join(base, f) for f in d["checkpoints"]]
        else:
            for pat in ("*.safetensors", "pytorch_model*.bin", "*.pt"):
                files = sorted(glob.glob(os.path.join(load_dir, pat)))
                if files:
                    break
        sd = {}
        for f in files:
            if f.endswith(".safetensors"):
                from safetensors.torch import load_file
                sd.update(load_file(f))
            else:
                part = torch.load(f, map_location="cpu", weights_only=False)
                sd.update(part.get("module", part.get("model", part)) if isinstance(part, dict) else part)
        if not sd:
            raise ValueError(f"no checkpoint files found for {load_dir}")
        # ---- a skeleton built on the meta device (``OnDevice(device="meta")``): tensor-parallel layers could not slice
        # anything when they were created, so each one is materialised from the checkpoint and sliced by its own rule
        # right away (one layer whole at a time), then everything else is assigned
        from ..module_inject.layers import TensorParallel_Layer
        done = set()
        for name, mod in self.module.named_modules():
            w = getattr(mod, "weight", None)
            if not isinstance(mod, TensorParallel_Layer) or w is None or not w.is_meta:
                continue
            wk, bk = f"{name}.weight", f"{name}.bias"
            if wk not in sd:
                continue
            mod.weight = nn.Parameter(sd[wk].to(w.dtype), requires_grad=False)
            b = getattr(mod, "bias", None)
            if b is not None and bk in sd:
                mod.bias = nn.Parameter(sd[bk].to(b.dtype), requires_grad=False)
            mod._tp_partition([mod.weight, getattr(mod, "bias", None)])
            done |= {wk, bk}
        own = self.module.state_dict()
        tp_rank = dist.get_rank(group=self.mp_group) if self.mp_group is not None else 0
        tp = self._config.tensor_parallel.tp_size
        fixed, on_meta = {}, False
        for k, v in sd.items():
            if k not in own or k in done:
                continue
            tgt = own[k]
            on_meta = on_meta or tgt.is_meta
            if v.shape == tgt.shape:
                fixed[k] = v.to(tgt.dtype) if tgt.is_meta else v
                continue
            for dim in range(v.dim()):          # full tensor -> this rank's TP slice
                if v.shape[dim] == tgt.shape[dim] * tp:
                    fixed[k] = v.chunk(tp, dim=dim)[tp_rank]
                    break
        missing = self.module.load_state_dict(fixed, strict=False, assign=on_meta)
        left = [k for k in missing.missing_keys if k not in done]
        if load_module_strict and left:
            logger.warning(f"checkpoint is missing {len(left)} keys, e.g. {left[:3]}")
        still_meta = [n for n, p in self.module.named_parameters() if p.is_meta]
        if still_meta:
            tied = getattr(self.module, "tie_weights", None)
            if callable(tied):
                tied()                           # e.g. lm_head sharing the embedding: re-tie to the loaded tensor
                still_meta = [n for n, p in self.module.named_parameters() if p.is_meta]
        if still_meta:
            raise RuntimeError(f"{len(still_meta)} parameters have no data after loading the checkpoint "
                               f"(e.g. {still_meta[:3]}): the checkpoint does not cover the model")
        self._materialise_meta_buffers(sd)

    @staticmethod
    def _rebuilt_on_cpu(mod):
        """A parameter-free module whose buffers are computed in its constructor from a config object (rotary tables):
        build a second instance on the host to obtain them."""
        if any(True for _ in mod.parameters(recurse=False)) or not hasattr(mod, "config"):
            return None
        try:
            with torch.device("cpu"):
                return type(mod)(mod.config)
        except Exception:
            return None

    def _materialise_meta_buffers(self, sd) -> None:
        """Buffers of a meta-device skeleton: persistent ones come from the checkpoint; the usual non-persistent one —
        the rotary ``inv_freq`` table — is recomputed by constructing the (parameter-free) module again on the host; anything else is an error here
        rather than a failure inside ``.to(device)`` later."""
        lost = []
        for name, mod in self.module.named_modules():
            for bname, buf in list(mod._buffers.items()):
                if buf is None or not buf.is_meta:
                    continue
                key = f"{name}.{bname}" if name else bname
                if key in sd:
                    mod._buffers[bname] = sd[key].to(buf.dtype)
                else:
                    fresh = self._rebuilt_on_cpu(mod)
                    if fresh is not None and isinstance(fresh._buffers.get(bname), torch.Tensor):
                        mod._buffers[bname] = fresh._buffers[bname].to(buf.dtype)
                    else:
                        lost.append(key)
        if lost:
            raise RuntimeError(f"buffers without data after loading the checkpoint: {lost[:5]} — build the skeleton "
                               "with these buffers on a real device or add them to the checkpoint")

    # ------------------------------------------------------------- tensor-parallel sharded checkpoints
    _TP_PARTITIONS = 8       # files per rank, so that no single file holds all layers (reference: 8)

    def _injected_prefixes(self):
        from ..model_implementations.transformers.ds_transformer import DeepSpeedTransformerInference
        return tuple(n + "." for n, m in self.module.named_modules() if isinstance(m, DeepSpeedTransformerInference))

    def _save_mp_checkpoint(self, path: str) -> None:
        """Write the model AS IT IS NOW — injected, tensor-parallel sliced, possibly int8 — so that the next start
        loads each rank's slice directly instead of re-reading and re-slicing the full checkpoint (reference
        ``replace_module.py:482-545``): ``non-tp.pt`` (everything outside the injected layers, written once),
        ``tp_<rank>_<part>.pt`` (this rank's injected-layer tensors; an int8 weight is stored as ``[codes, scale]``) and
        the descriptor ``ds_inference_config.json`` that ``init_inference(checkpoint=...)`` takes."""
        import json
        from collections import OrderedDict
        os.makedirs(path, exist_ok=True)
        tp = self._config.tensor_parallel.tp_size
        rank = dist.get_rank(group=self.mp_group) if (self.mp_group is not None and tp > 1) else 0
        writer = (not dist.is_initialized()) or dist.get_rank() == 0
        prefixes = self._injected_prefixes()
        sd = self.module.state_dict()
        for name, p in self.module.named_parameters():
            if hasattr(p, "scale") and p.dtype == torch.int8:
                sd[name] = [p.data, p.scale]
        sliced = [k for k in sd if k.startswith(prefixes)] if prefixes else []
        if writer:
            torch.save(OrderedDict((k, v) for k, v in sd.items() if k not in set(sliced)), os.path.join(path, "non-tp.pt"))
            kind = type(self.module).__name__
            reprs = {torch.float32: "float32", torch.float16: "float16", torch.bfloat16: "bfloat16", torch.int8: "int8"}
            with open(os.path.join(path, "ds_inference_config.json"), "w") as f:
                json.dump({"type": "bloom" if "bloom" in kind.lower() else "ds_model", "base_dir": path,
                           "checkpoints": {"non_tp": ["non-tp.pt"],
                                           "tp": [f"tp_{r:0>2d}_{m:0>2d}.pt" for m in range(self._TP_PARTITIONS)
                                                  for r in range(tp)]},
                           "version": 1.0, "parallelization": "tp", "tp_size": tp,
                           "dtype": reprs.get(self._config.dtype, str(self._config.dtype))}, f)
        per = len(sliced) // self._TP_PARTITIONS + 1
        for m in range(self._TP_PARTITIONS):
            part = OrderedDict((k, sd[k]) for k in sliced[m * per:(m + 1) * per])
            torch.save(part, os.path.join(path, f"tp_{rank:0>2d}_{m:0>2d}.pt"))
        if dist.is_initialized():
            dist.barrier()

    def _load_mp_checkpoint(self, desc: dict) -> None:
        """Inverse of :meth:`_save_mp_checkpoint`: tensors already have this rank's shapes."""
        tp = self._config.tensor_parallel.tp_size
        if int(desc.get("tp_size", 1)) != tp:
            raise ValueError(f"the sharded checkpoint was written for tp_size={desc.get('tp_size')}, "
                             f"this engine runs tp_size={tp}; load the original checkpoint instead")
        base = desc.get("base_dir", self._config.base_dir)
        rank = dist.get_rank(group=self.mp_group) if (self.mp_group is not None and tp > 1) else 0
        files = list(desc["checkpoints"]["non_tp"]) + [f for f in desc["checkpoints"]["tp"]
                                                       if f.startswith(f"tp_{rank:0>2d}_")]
        params = dict(self.module.named_parameters())
        plain = {}
        for f in files:
            for k, v in torch.load(os.path.join(base, f), map_location="cpu", weights_only=False).items():
                if isinstance(v, (list, tuple)):                      # int8 codes + per-group step sizes
                    codes, scale = v
                    owner, _, leaf = k.rpartition(".")
                    mod = self.module.get_submodule(owner) if owner else self.module
                    new = nn.Parameter(codes.to(params[k].device) if k in params else codes, requires_grad=False)
                    new.scale = scale
                    setattr(mod, leaf, new)
                else:
                    plain[k] = v
        missing = self.module.load_state_dict(plain, strict=False)
        left = [k for k in missing.missing_keys if not (k in params and params[k].dtype == torch.int8)]
        if left:
            logger.warning(f"sharded checkpoint is missing {len(left)} keys, e.g. {left[:3]}")

    # ------------------------------------------------------------------------------ CUDA graphs
    def _create_cuda_graph(self, *inputs, **kwargs):
        stream = torch.cuda.Stream()
        stream.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(stream):
            for _ in range(3):
                self.module(*inputs, **kwargs)
        torch.cuda.current_stream().wait_stream(stream)
        self._cuda_graphs = torch.cuda.CUDAGraph()
        self.static_inputs, self.static_kwargs = inputs, kwargs
        with torch.cuda.graph(self._cuda_graphs):
            self.static_output = self.module(*self.static_inputs, **self.static_kwargs)
        self.cuda_graph_created = True

    def _graph_replay(self, *inputs, **kwargs):
        for i, t in enumerate(inputs):
            if torch.is_tensor(t):
                self.static_inputs[i].copy_(t)
        for k, t in kwargs.items():
            if torch.is_tensor(t):
                self.static_kwargs[k].copy_(t)
        self._cuda_graphs.replay()
        return self.static_output

    # ------------------------------------------------------------------------------------ calls
    def forward(self, *inputs, **kwargs):
        start = None
        if self.model_profile_enabled and self._config.enable_cuda_graph:
            torch.cuda.synchronize()
            start = time.time()
        if self._config.enable_cuda_graph and not self.local_cuda_graph:
            if self.cuda_graph_created:
                outputs = self._graph_replay(*inputs, **kwargs)
            else:
                self._create_cuda_graph(*inputs, **kwargs)
                outputs = self._graph_replay(*inputs, **kwargs)
        else:
            outputs = self.module(*inputs, **kwargs)
        if start is not None:
            torch.cuda.synchronize()
            self._model_times.append((time.time() - start) * 1e3)
        return outputs

    def destroy(self):
        """Release engine-owned state: KV workspace of the injected layers, CUDA graphs, the shared TP group handle."""
        from ..model_implementations.transformers.ds_transformer import DeepSpeedTransformerInference
        DeepSpeedTransformerInference.layer_id = 0
        if DeepSpeedTransformerInference.workspace is not None:
            DeepSpeedTransformerInference.workspace.release_workspace()
            DeepSpeedTransformerInference.workspace = None
        self._cuda_graphs = None
        self.cuda_graph_created = False
        InferenceEngine.inference_mp_group = None

    def is_compiled(self) -> bool:
        return getattr(self, "_is_compiled", False)

    def load_model_with_checkpoint(self, r_module):
        """Stream the configured checkpoint's state dicts into ``r_module`` (meta-device models are materialised)."""
        from ..module_inject.load_checkpoint import load_model_with_checkpoint
        sd = getattr(self, "sd", None)
        if sd is None:
            raise RuntimeError("no checkpoint state dict is loaded: pass `checkpoint=` to init_inference")
        return load_model_with_checkpoint(r_module, sd, getattr(self, "mp_replace", None),
                                          ckpt_type=getattr(self, "ckpt_type", "pp"),
                                          ckpt_mp_size=getattr(self, "ckpt_mp_size", 1), container=None)

    def build_alibi_tensor(self):
        """Tensor-parallel ranks hold a slice of the heads: patch the client's ALiBi builder so each rank gets ITS
        heads' slopes (BLOOM-style models; reference ``inference/engine.py:197-206``)."""
        from ..module_inject import auto_tp_model_utils as u
        tr = getattr(self.module, "transformer", None)
        if tr is None:
            return
        if hasattr(tr, "build_alibi_tensor"):
            tr.build_alibi_tensor = u.build_bloom_alibi_tensor
        if hasattr(tr, "build_mpt_alibi_tensor"):
            from types import MethodType
            tr.build_mpt_alibi_tensor_orig = tr.build_mpt_alibi_tensor
            tr.build_mpt_alibi_tensor = MethodType(u.build_mpt_alibi_tensor, tr)
        model = getattr(self.module, "model", None)
        if model is not None and hasattr(model, "get_alibi_mask"):
            from types import MethodType
            model.get_alibi_mask_orig = model.get_alibi_mask
            model.get_alibi_mask = MethodType(u.get_alibi_mask, model)

    def build_attn_bias(self):
        tr = getattr(self.module, "transformer", None)
        if tr is not None and hasattr(tr, "_attn_bias"):
            from types import MethodType
            from ..module_inject.auto_tp_model_utils import build_mpt_atten_bias_tensor
            tr._attn_bias_orig = tr._attn_bias
            tr._attn_bias = MethodType(build_mpt_atten_bias_tensor, tr)

    def remove_mask_prepare_for_bloom(self):
        """The fused BLOOM attention builds its own causal mask: make the client's ``_prepare_attn_mask`` a
        pass-through."""
        tr = getattr(self.module, "transformer", None)
        if tr is not None and hasattr(tr, "_prepare_attn_mask"):
            tr._prepare_attn_mask = lambda attention_mask, *args, **kwargs: attention_mask

    def _generate(self, *inputs, **kwargs):
        num_beams = kwargs.get("num_beams", 1)
        if "generation_config" in kwargs:
            num_beams = getattr(kwargs["generation_config"], "num_beams", num_beams)
        if num_beams > 1:
            raise NotImplementedError("DeepSpeed does not support `num_beams` > 1, if this is important to you please "
                                      "add your request to: https://github.com/deepspeedai/DeepSpeed/issues/2506")
        if "input_ids" in kwargs and self._config.max_out_tokens:
            n = kwargs["input_ids"].shape[-1]
            if n > self._config.max_out_tokens:
                raise RuntimeError(f"Input with size {n} exceeds maximum length of {self._config.max_out_tokens}. "
                                   "Please increase `max_tokens` in the DeepSpeed Inference Config.")
        return self.module.generate(*inputs, **kwargs)
