import math

import torch

from shuffle_exchange_b200.runtime import lr_schedules as L


def _opt(lr=0.1):
    return torch.optim.SGD([torch.nn.Parameter(torch.zeros(1))], lr=lr)


def test_warmup_lr_log_and_linear():
    o = _opt()
    s = L.WarmupLR(o, warmup_min_lr=0.0, warmup_max_lr=1.0, warmup_num_steps=10, warmup_type="linear")
    vals = []
    for _ in range(15):
        s.step()
        vals.append(o.param_groups[0]["lr"])
    assert abs(vals[4] - 0.4) < 1e-9 and vals[-1] == 1.0
    o = _opt()
    s = L.WarmupLR(o, warmup_max_lr=1.0, warmup_num_steps=10, warmup_type="log")
    s.step(); s.step()
    assert abs(o.param_groups[0]["lr"] - math.log(2) / math.log(10)) < 1e-9


def test_warmup_decay_reaches_zero():
    o = _opt()
    s = L.WarmupDecayLR(o, total_num_steps=20, warmup_max_lr=1.0, warmup_num_steps=5, warmup_type="linear")
    for _ in range(21):
        s.step()
    assert o.param_groups[0]["lr"] == 0.0


def test_cosine_and_state_dict():
    o = _opt(lr=2.0)
    s = L.WarmupCosineLR(o, total_num_steps=100, warmup_num_steps=10, cos_min_ratio=0.1, warmup_type="linear")
    for _ in range(100):
        s.step()
    assert abs(o.param_groups[0]["lr"] - 0.2) < 1e-2
    sd = s.state_dict()
    s2 = L.WarmupCosineLR(_opt(lr=2.0), total_num_steps=100, warmup_num_steps=10, cos_min_ratio=0.1)
    s2.load_state_dict(sd)
    assert s2.last_batch_iteration == s.last_batch_iteration


def test_one_cycle_and_range_test():
    o = torch.optim.Adam([torch.nn.Parameter(torch.zeros(1))], lr=0.1)
    s = L.OneCycle(o, cycle_min_lr=0.1, cycle_max_lr=1.0, cycle_first_step_size=10, decay_step_size=5, decay_lr_rate=0.5)
    lrs = []
    for _ in range(40):
        s.step()
        lrs.append(o.param_groups[0]["lr"])
    assert max(lrs) <= 1.0 + 1e-9 and abs(lrs[9] - 1.0) < 1e-9 and lrs[-1] < 0.1
    o = _opt()
    r = L.LRRangeTest(o, lr_range_test_min_lr=0.01, lr_range_test_step_size=5, lr_range_test_step_rate=2.0,
                      lr_range_test_staircase=True)
    for _ in range(10):
        r.step()
    assert abs(o.param_groups[0]["lr"] - 0.01 * (1 + 2.0 * 2)) < 1e-9
