"""DeepSpeedCPUAdagrad (reference: deepspeed/ops/adagrad/cpu_adagrad.py:12-126; native
create/update/destroy :22-29, :90-96) on the host kernel in csrc/host/cpu_optim.cpp."""
import torch

from ...native import host as _h


class DeepSpeedCPUAdagrad(torch.optim.Optimizer):
    optimizer_id = 0

    def __init__(self, model_params, lr=1e-2, eps=1e-10, weight_decay=0, amsgrad=False, fp32_optimizer_states=True):
        super().__init__(model_params, dict(lr=lr, eps=eps, weight_decay=weight_decay, amsgrad=amsgrad))
        self.opt_id = DeepSpeedCPUAdagrad.optimizer_id
        DeepSpeedCPUAdagrad.optimizer_id += 1
        self.fp32_optimizer_states = fp32_optimizer_states
        self._lib = _h.load()
        self._lib.sxbh_create_adagrad(self.opt_id, lr, eps, weight_decay, 0)

    def __del__(self):
        try:
            self._lib.sxbh_destroy_adagrad(self.opt_id)
        except Exception:
            pass

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        for group in self.param_groups:
            for p in group["params"]:
                if p.grad is None:
                    continue
                assert p.device.type == "cpu", "CPUAdagrad expects host parameters"
                st = self.state[p]
                if len(st) == 0:
                    st["step"] = 0
                    st["exp_avg_sq"] = torch.zeros_like(p.data, dtype=torch.float32, device="cpu")
                st["step"] += 1
                g = p.grad.data
                if g.is_sparse:
                    g = g.to_dense()
                g = g.float().contiguous()
                self._lib.sxbh_adagrad_update(self.opt_id, st["step"], float(group["lr"]), float(group["eps"]),
                                              float(group["weight_decay"]), p.data.data_ptr(), g.data_ptr(),
                                              st["exp_avg_sq"].data_ptr(), p.numel(), None, 0)
        return loss
