"""Per-block dominant Hessian eigenvalue by power iteration on Hessian-vector products.

Reference: deepspeed/runtime/eigenvalue.py:13-160 — for each of ``layer_num`` blocks under ``layer_name``,
iterate ``v <- normalize(H v)`` (``H v`` via double backward through ``param.grad``'s graph, so the engine must
have run ``backward(create_graph=True)``), stop on relative change < ``tol`` or ``max_iter``; results are
normalised to [0, 1] by the largest magnitude and returned as ``{id(param): (eigenvalue, block)}``.
Used by the MoQ quantisation schedule (reference engine.py:2472-2491).
"""
from __future__ import annotations

import torch

from ..utils.logging import log_dist


class Eigenvalue:
    def __init__(self, verbose=False, max_iter=100, tol=1e-2, stability=0, gas_boundary_resolution=1,
                 layer_name="", layer_num=0):
        self.verbose, self.max_iter, self.tol, self.stability = verbose, max_iter, tol, stability
        self.gas_boundary_resolution = gas_boundary_resolution
        self.layer_name, self.layer_num = layer_name, layer_num
        assert len(layer_name) > 0 and layer_num > 0
        log_dist(f"enabled eigenvalue with verbose={verbose}, max_iter={max_iter}, tol={tol}, stability={stability}, "
                 f"gas_boundary_resolution={gas_boundary_resolution}, layer_name={layer_name}, layer_num={layer_num}",
                 ranks=[0])

    # -- small vector-list algebra ----------------------------------------------------------------
    @staticmethod
    def nan_to_num(x: torch.Tensor) -> torch.Tensor:
        return torch.nan_to_num(x, nan=0.0, posinf=0.0, neginf=0.0)

    @staticmethod
    def inner_product(xs, ys) -> torch.Tensor:
        return sum(torch.sum(x * y) for x, y in zip(xs, ys))

    def normalize(self, vs):
        norm = self.inner_product(vs, vs) ** 0.5 + self.stability
        return [self.nan_to_num(v / norm) for v in vs]

    def get_layers(self, module):
        m = module
        for name in self.layer_name.split("."):
            assert hasattr(m, name), "layer_name configuration is invalid."
            m = getattr(m, name)
        return m

    # -- main entry --------------------------------------------------------------------------------
    def compute_eigenvalue(self, module, device=None, scale=1.0):
        layers = self.get_layers(module)
        values, keys = [], []
        for b in range(self.layer_num):
            block = layers[b]
            live = [p for p in block.parameters() if p.grad is not None and p.grad.grad_fn is not None]
            keys.append([id(p) for p in block.parameters()])
            if not live:
                log_dist("The model does NOT support eigenvalue computation.", ranks=[0])
                return []
            grads = [p.grad for p in live]
            rng = torch.random.get_rng_state()           # probing must not perturb the training RNG stream
            v = [torch.randn(p.size(), device=device if device is not None else p.device) for p in live]
            torch.random.set_rng_state(rng)
            v = self.normalize(v)
            cur, prev, it = 1.0, 0.0, 0
            while it < self.max_iter and abs(cur) > 0 and abs((cur - prev) / cur) >= self.tol:
                prev = cur
                hv = torch.autograd.grad(grads, live, grad_outputs=v, only_inputs=True, retain_graph=True)
                hv = [self.nan_to_num(h).float() for h in hv]
                cur = float(self.inner_product(hv, v))
                v = [x / scale for x in self.normalize(hv)]
                it += 1
            cur *= scale
            values.append(cur)
            if self.verbose:
                log_dist(f"block: {b}, power iteration: {it}, eigenvalue: {cur}", ranks=[0])
        values = self.post_process(values)
        out = {}
        for i, (ks, val) in enumerate(zip(keys, values)):
            out.update(dict.fromkeys(ks, (val, i)))
        return out

    @staticmethod
    def post_process(values):
        top = abs(max(values, key=abs)) if values else 1.0
        return [abs(v) / top if v != 0.0 else 1.0 for v in values]
