"""DeepSpeedCPULion (reference: deepspeed/ops/lion/cpu_lion.py:13-141; native create/update/destroy
:57-64, :125)."""
import torch

from ...native import host as _h


class DeepSpeedCPULion(torch.optim.Optimizer):
    optimizer_id = 0

    def __init__(self, model_params, lr=1e-3, betas=(0.9, 0.999), weight_decay=0, fp32_optimizer_states=True):
        super().__init__(model_params, dict(lr=lr, betas=betas, weight_decay=weight_decay))
        self.opt_id = DeepSpeedCPULion.optimizer_id
        DeepSpeedCPULion.optimizer_id += 1
        self.fp32_optimizer_states = fp32_optimizer_states
        self._lib = _h.load()
        self._lib.sxbh_create_lion(self.opt_id, lr, betas[0], betas[1], weight_decay, 0)

    def __del__(self):
        try:
            self._lib.sxbh_destroy_lion(self.opt_id)
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
                assert p.device.type == "cpu", "CPULion expects host parameters"
                st = self.state[p]
                if len(st) == 0:
                    st["step"] = 0
                    st["exp_avg"] = torch.zeros_like(p.data, dtype=torch.float32, device="cpu")
                st["step"] += 1
                b1, b2 = group["betas"]
                g = p.grad.data.float().contiguous()
                self._lib.sxbh_lion_update(self.opt_id, st["step"], float(group["lr"]), float(b1), float(b2),
                                           float(group["weight_decay"]), p.data.data_ptr(), g.data_ptr(),
                                           st["exp_avg"].data_ptr(), p.numel(), None, 0)
        return loss
