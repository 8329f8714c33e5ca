import torch

from shuffle_exchange_b200.runtime.fp16.loss_scaler import CreateLossScaler, DynamicLossScaler, LossScaler


def test_dynamic_scaler_policy():
    s = CreateLossScaler(torch.half, 0, True, {"init_scale": 2 ** 8, "scale_window": 2, "delayed_shift": 2,
                                               "min_scale": 1, "consecutive_hysteresis": False})
    assert isinstance(s, DynamicLossScaler) and s.cur_scale == 256
    s.update_scale(True)            # hysteresis absorbs the first overflow
    assert s.cur_scale == 256
    s.update_scale(True)
    assert s.cur_scale == 128
    s.update_scale(False); s.update_scale(False)
    assert s.cur_scale == 256
    st = CreateLossScaler(torch.bfloat16, 128.0, False, None)
    assert isinstance(st, LossScaler) and st.cur_scale == 1.0
