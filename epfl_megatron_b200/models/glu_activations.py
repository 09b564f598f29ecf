"""GLU activation modules (megatron/model/glu_activations.py:11-54 surface):
``x1, x2 = chunk(x, 2); x1 * act(x2)`` -- the second half is the gate."""
import torch
from torch import nn

from ..ops import functional as Fn


class _GLUBaseModule(nn.Module):
    kind = None

    def forward(self, x):
        return Fn.glu(x, self.kind)


class LiGLU(_GLUBaseModule):
    kind = "liglu"


class GEGLU(_GLUBaseModule):
    kind = "geglu"


class ReGLU(_GLUBaseModule):
    kind = "reglu"


class SwiGLU(_GLUBaseModule):
    kind = "swiglu"


liglu, geglu, reglu, swiglu = LiGLU(), GEGLU(), ReGLU(), SwiGLU()

GLU_ACTIVATIONS = {"geglu": geglu, "liglu": liglu, "reglu": reglu, "swiglu": swiglu}
