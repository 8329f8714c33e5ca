"""Stream parameters out of a local HuggingFace checkpoint directory (reference ``checkpoint/huggingface_engine.py``:
the reference also downloads from the hub; there is no network here, so ``model_name_or_path`` must be a directory).
Understands ``*.safetensors`` (single or sharded with an index json) and ``pytorch_model*.bin``."""
from __future__ import annotations

import json
import os
from typing import Iterable, List, Tuple

import torch

from ..logging import inference_logger
from .base_engine import CheckpointEngineBase


class HuggingFaceCheckpointEngine(CheckpointEngineBase):

    def __init__(self, model_name_or_path: str, auth_token: str = None, **hf_kwargs) -> None:
        super().__init__()
        self.model_name_or_path = model_name_or_path
        self.auth_token = auth_token
        if not os.path.isdir(model_name_or_path):
            raise ValueError(f"{model_name_or_path} is not a local directory; hub downloads are unavailable offline")
        from transformers import AutoConfig, GenerationConfig
        self.model_config = AutoConfig.from_pretrained(model_name_or_path, **hf_kwargs)
        try:
            self.generation_config = GenerationConfig.from_pretrained(model_name_or_path)
        except Exception:
            self.generation_config = None
        if not hasattr(self.model_config, "max_seq_length"):
            for k in ("max_position_embeddings", "n_positions", "seq_length"):
                if hasattr(self.model_config, k):
                    self.model_config.max_seq_length = getattr(self.model_config, k)
                    break
            else:
                self.model_config.max_seq_length = 2048
        self._local_checkpoint_dir = model_name_or_path
        self._all_ckpt_paths = self._fetch_checkpoint_files()

    def _fetch_checkpoint_files(self) -> List[str]:
        d = self._local_checkpoint_dir
        for index, single in (("model.safetensors.index.json", "model.safetensors"),
                              ("pytorch_model.bin.index.json", "pytorch_model.bin")):
            ip = os.path.join(d, index)
            if os.path.isfile(ip):
                with open(ip) as f:
                    files = sorted(set(json.load(f)["weight_map"].values()))
                return [os.path.join(d, x) for x in files]
            sp = os.path.join(d, single)
            if os.path.isfile(sp):
                return [sp]
        found = sorted(x for x in os.listdir(d) if x.endswith(".safetensors") or x.endswith(".bin"))
        if not found:
            raise FileNotFoundError(f"no *.safetensors / *.bin weights under {d}")
        return [os.path.join(d, x) for x in found]

    def parameters(self) -> Iterable[Tuple[str, torch.Tensor]]:
        for path in self._all_ckpt_paths:
            inference_logger().info(f"Loading checkpoint: {path}")
            if path.endswith(".safetensors"):
                from safetensors import safe_open
                with safe_open(path, framework="pt", device="cpu") as f:
                    for name in f.keys():
                        yield name, f.get_tensor(name)
            else:
                sd = torch.load(path, map_location="cpu", weights_only=True)
                for name, t in sd.items():
                    yield name, t
                del sd
