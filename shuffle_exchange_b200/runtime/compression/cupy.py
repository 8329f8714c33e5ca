"""``CupyBackend`` (reference: ``deepspeed/runtime/compression/cupy.py``): the reference converts torch tensors to cupy
through DLPack to use ``cupy.packbits``; here the same interface runs on the framework's ``sxb_packbits`` /
``sxb_unpackbits`` kernels, so "cupy arrays" are simply torch tensors and the conversions are identities."""
from __future__ import annotations

from typing import List

import torch

from ..comm.compressed import packbits, unpackbits


class CupyBackend:

    def torch2cupy(self, tensor: torch.Tensor) -> torch.Tensor:
        return tensor

    def cupy2torch(self, cupy_tensor: torch.Tensor) -> torch.Tensor:
        return cupy_tensor

    def compress_by_chunk(self, cupy_bool_tensor: torch.Tensor, num_chunks: int) -> List[torch.Tensor]:
        """Pack a boolean (sign >= 0) tensor into bits, one packed tensor per equal chunk."""
        signed = cupy_bool_tensor.to(torch.float32) * 2 - 1 if cupy_bool_tensor.dtype == torch.bool else cupy_bool_tensor
        return [packbits(c.contiguous()) for c in signed.reshape(-1).chunk(num_chunks)]

    def decompress(self, packed: torch.Tensor, n: int, dtype=torch.float32, scale: float = 1.0) -> torch.Tensor:
        return unpackbits(packed, n, dtype, scale)
