"""Error-compensated 1-bit all-reduce (the transport of OnebitAdam / OnebitLamb / ZeroOneAdam).

Reference: ``deepspeed/runtime/comm/compressed.py:57-137`` (and the equivalent ``nccl.py`` / ``mpi.py`` / ``hccl.py``
variants, which differ only in how bits are packed).  Algorithm, per call on a flat momentum buffer of ``n`` elements
(padded to a multiple of ``8 * world``):

1. worker: ``x += worker_error``; ``scale = ||x||_2 / sqrt(n)``; ``worker_error = x - scale * sign(x)``; pack the
   signs to bits (``sxb_packbits``);
2. all-to-all the sign chunks (rank r receives chunk r of everyone), all-gather the worker scales;
3. server (each rank for its chunk): average ``scale_r * sign_r``; add ``server_error``; again take scale + sign and
   update ``server_error``;
4. all-gather the packed server signs and server scales; unpack into the result.

Traffic per rank: ``n/8`` bytes out + ``n/8`` bytes in instead of ``2 n * sizeof(dtype)``.
"""
from __future__ import annotations

import torch

from ... import comm as dist
from ...native import lib as _lib

_lib.register({"sxb_packbits": "pp" "l" "i" "p", "sxb_unpackbits": "pp" "l" "f" "i" "p"})


def packbits(x: torch.Tensor) -> torch.Tensor:
    """Sign bits of ``x`` (>= 0 -> 1), most significant bit first, ``ceil(n/8)`` bytes."""
    n = x.numel()
    out = torch.empty((n + 7) // 8, dtype=torch.uint8, device=x.device)
    if x.is_cuda:
        _lib.call("sxb_packbits", x.data_ptr(), out.data_ptr(), n, _lib.dtype_code(x.dtype), _lib.stream_ptr())
        return out
    bits = (x.reshape(-1) >= 0).to(torch.uint8)
    pad = (-n) % 8
    if pad:
        bits = torch.cat([bits, bits.new_zeros(pad)])
    weights = torch.tensor([128, 64, 32, 16, 8, 4, 2, 1], dtype=torch.uint8)
    return (bits.view(-1, 8) * weights).sum(dim=1).to(torch.uint8)


def unpackbits(packed: torch.Tensor, n: int, dtype=torch.float32, scale: float = 1.0) -> torch.Tensor:
    """Inverse of :func:`packbits`: ``+scale`` where the bit is set, ``-scale`` elsewhere."""
    out = torch.empty(n, dtype=dtype, device=packed.device)
    if packed.is_cuda:
        _lib.call("sxb_unpackbits", packed.data_ptr(), out.data_ptr(), n, float(scale), _lib.dtype_code(dtype),
                  _lib.stream_ptr())
        return out
    shifts = torch.tensor([7, 6, 5, 4, 3, 2, 1, 0], dtype=torch.uint8)
    bits = ((packed.view(-1, 1) >> shifts) & 1).reshape(-1)[:n]
    return torch.where(bits.bool(), torch.tensor(scale), torch.tensor(-scale)).to(dtype)


class CompressedBackend:
    def __init__(self, mpu=None):
        self.world_group = dist.new_group(ranks=list(range(dist.get_world_size()))) if mpu is None else mpu.get_data_parallel_group()
        self.size = dist.get_world_size(group=self.world_group)
        self.rank = dist.get_rank(group=self.world_group)

    def my_igather(self, rank, size, group, sendbuf, recvbuf, root):
        req = []
        if rank == root:
            for idx in range(size):
                if idx != rank:
                    req.append(dist.irecv(recvbuf[idx], src=idx, group=group))
                else:
                    recvbuf[rank] = sendbuf
        else:
            req.append(dist.isend(sendbuf, group=group, dst=root))
        return req

    def my_gather(self, rank, size, group, sendbuf, recvbuf, root):
        for r in self.my_igather(rank, size, group, sendbuf, recvbuf, root):
            r.wait()

    @torch.no_grad()
    def compressed_allreduce(self, buffer_m: torch.Tensor, worker_error: torch.Tensor, server_error: torch.Tensor,
                             local_rank: int = 0) -> torch.Tensor:
        original_shape, original_size = buffer_m.size(), buffer_m.numel()
        world = self.size
        if original_size != worker_error.numel():
            flat = torch.zeros(worker_error.numel(), dtype=buffer_m.dtype, device=buffer_m.device)
            flat[:original_size] = buffer_m.reshape(-1)
            buffer_m = flat
        else:
            buffer_m = buffer_m.reshape(-1)
        n = buffer_m.numel()
        assert n % (8 * world) == 0, "error-feedback buffers must be padded to a multiple of 8 * world_size"
        chunk = n // world
        # ---- worker side
        buffer_m.add_(worker_error)
        worker_scale = torch.linalg.vector_norm(buffer_m.float()) / (n ** 0.5)
        packed = packbits(buffer_m)
        worker_error.copy_(buffer_m - unpackbits(packed, n, buffer_m.dtype) * worker_scale.to(buffer_m.dtype))
        # ---- exchange: signs by all-to-all, scales by all-gather
        recv_signs = torch.empty_like(packed)
        dist.all_to_all_single(recv_signs, packed, group=self.world_group)
        scales = [torch.empty_like(worker_scale.reshape(1)) for _ in range(world)]
        dist.all_gather(scales, worker_scale.reshape(1), group=self.world_group)
        # ---- server side (this rank owns chunk `rank`)
        server_m = torch.zeros(chunk, dtype=torch.float32, device=buffer_m.device)
        per = chunk // 8
        for r in range(world):
            server_m.add_(unpackbits(recv_signs[r * per: (r + 1) * per], chunk, torch.float32) * scales[r].float())
        server_m.div_(world).add_(server_error.float())
        server_scale = torch.linalg.vector_norm(server_m) / (chunk ** 0.5)
        server_packed = packbits(server_m)
        server_error.copy_((server_m - unpackbits(server_packed, chunk, torch.float32) * server_scale).to(server_error.dtype))
        # ---- broadcast the server result
        all_signs = torch.empty(world * per, dtype=torch.uint8, device=buffer_m.device)
        dist.all_gather_into_tensor(all_signs, server_packed, group=self.world_group)
        all_scales = [torch.empty_like(server_scale.reshape(1)) for _ in range(world)]
        dist.all_gather(all_scales, server_scale.reshape(1), group=self.world_group)
        result = unpackbits(all_signs, n, torch.float32).view(world, chunk)
        result.mul_(torch.cat(all_scales).float().view(world, 1))
        out = result.view(-1).to(buffer_m.dtype)
        if original_size != n:
            out = out[:original_size]
        return out.view(original_shape)
