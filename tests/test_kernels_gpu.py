"""GPU numerics: every CUDA kernel vs a plain PyTorch fp32 restatement of the op."""
import pytest

from epfl_megatron_b200.ops import selfcheck

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", list(selfcheck.CHECKS))
def test_kernel(name):
    selfcheck.CHECKS[name]()
