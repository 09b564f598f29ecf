"""Unpickling stubs for checkpoints written by very old Megatron versions.

Those checkpoints pickled their loss-scaler objects under ``megatron.fp16_deprecated.loss_scaler``
(megatron/fp16_deprecated/loss_scaler.py:1-26); ``torch.load`` needs classes of the same names and
attributes to rebuild them. Nothing here is used for training: the live scalers are in
``optimizer/grad_scaler.py``. ``megatron/__init__.py`` aliases this module under the old path.
"""


class LossScaler:
    def __init__(self, scale=1):
        self.cur_scale = scale


class DynamicLossScaler:
    def __init__(self, init_scale=2 ** 32, scale_factor=2.0, scale_window=1000, min_scale=1,
                 delayed_shift=1, consecutive_hysteresis=False):
        self.cur_scale = init_scale
        self.cur_iter = 0
        self.last_overflow_iter = -1
        self.scale_factor = scale_factor
        self.scale_window = scale_window
        self.min_scale = min_scale
        self.delayed_shift = delayed_shift
        self.cur_hysteresis = delayed_shift
        self.consecutive_hysteresis = consecutive_hysteresis
