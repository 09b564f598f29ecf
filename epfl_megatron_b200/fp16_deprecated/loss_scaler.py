"""Unpickling stand-ins for checkpoints written by very old Megatron versions.

Such checkpoints pickled their loss-scaler objects under ``megatron.fp16_deprecated.loss_scaler``
(megatron/fp16_deprecated/loss_scaler.py:1-26). Unpickling restores ``__dict__`` directly, so all
``torch.load`` needs is classes with these two names at that path (``megatron/__init__.py`` aliases this
module there); the constructors below only exist for code that builds one by hand. The scalers that
training actually uses are in ``optimizer/grad_scaler.py``.
"""


class _Restorable:
    _defaults = {}

    def __init__(self, **overrides):
        unknown = set(overrides) - set(self._defaults)
        if unknown:
            raise TypeError(f"unexpected arguments {sorted(unknown)}")
        self.__dict__.update({**self._defaults, **overrides})


class LossScaler(_Restorable):
    _defaults = {"cur_scale": 1}

    def __init__(self, scale=1):
        super().__init__(cur_scale=scale)


class DynamicLossScaler(_Restorable):
    _defaults = {"cur_scale": 2 ** 32, "cur_iter": 0, "last_overflow_iter": -1, "scale_factor": 2.0,
                 "scale_window": 1000, "min_scale": 1, "delayed_shift": 1, "cur_hysteresis": 1,
                 "consecutive_hysteresis": False}

    def __init__(self, init_scale=2 ** 32, scale_factor=2.0, scale_window=1000, min_scale=1, delayed_shift=1,
                 consecutive_hysteresis=False):
        super().__init__(cur_scale=init_scale, scale_factor=scale_factor, scale_window=scale_window,
                         min_scale=min_scale, delayed_shift=delayed_shift, cur_hysteresis=delayed_shift,
                         consecutive_hysteresis=consecutive_hysteresis)
