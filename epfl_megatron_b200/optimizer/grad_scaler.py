"""Loss scalers (megatron/optimizer/grad_scaler.py:11-120 behaviour): a constant scale, or
a dynamic one that backs off after ``hysteresis`` overflowing steps and grows after
``growth_interval`` clean ones."""
from abc import ABC, abstractmethod

import torch


def _dev():
    return torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu")


class MegatronGradScaler(ABC):
    def __init__(self, initial_scale):
        assert initial_scale > 0.0
        self._scale = torch.tensor([initial_scale], dtype=torch.float, device=_dev())

    @property
    def scale(self):
        return self._scale

    @property
    def inv_scale(self):
        return self._scale.double().reciprocal().float()

    @abstractmethod
    def update(self, found_inf):
        pass

    @abstractmethod
    def state_dict(self):
        pass

    @abstractmethod
    def load_state_dict(self, state_dict):
        pass


class ConstantGradScaler(MegatronGradScaler):
    def update(self, found_inf):
        pass

    def state_dict(self):
        return dict()

    def load_state_dict(self, state_dict):
        pass


class DynamicGradScaler(MegatronGradScaler):
    def __init__(self, initial_scale, min_scale, growth_factor, backoff_factor, growth_interval, hysteresis):
        super().__init__(initial_scale)
        assert 0.0 < min_scale <= initial_scale
        assert growth_factor > 1.0 and 0.0 < backoff_factor < 1.0
        assert growth_interval > 0 and hysteresis > 0
        self.min_scale = torch.tensor([min_scale], dtype=torch.float, device=_dev())
        self.growth_factor = torch.tensor([growth_factor], dtype=torch.float, device=_dev())
        self.backoff_factor = torch.tensor([backoff_factor], dtype=torch.float, device=_dev())
        self.growth_interval, self.hysteresis = growth_interval, hysteresis
        self._growth_tracker = 0
        self._hysteresis_tracker = hysteresis

    def update(self, found_inf):
        if found_inf:
            self._growth_tracker = 0
            self._hysteresis_tracker -= 1
            if self._hysteresis_tracker <= 0:
                self._scale = torch.max(self._scale * self.backoff_factor, self.min_scale)
        else:
            self._growth_tracker += 1
            if self._growth_tracker == self.growth_interval:
                self._growth_tracker = 0
                self._hysteresis_tracker = self.hysteresis
                self._scale = self._scale * self.growth_factor

    def state_dict(self):
        return {"scale": self._scale, "growth_tracker": self._growth_tracker,
                "hysteresis_tracker": self._hysteresis_tracker}

    def load_state_dict(self, state_dict):
        self._scale = state_dict["scale"].to(_dev())
        self._growth_tracker = state_dict["growth_tracker"]
        self._hysteresis_tracker = state_dict["hysteresis_tracker"]
