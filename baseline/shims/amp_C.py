"""torch._foreach_* stand-ins for the two amp_C kernels the reference calls."""
import torch


def multi_tensor_l2norm(noop_flag, tensor_lists, per_tensor):
    ts = tensor_lists[0]
    if not ts:
        z = torch.zeros(1, device="cuda")
        return z, z
    norms = torch._foreach_norm(ts)
    stacked = torch.stack([n.float() for n in norms])
    total = torch.linalg.vector_norm(stacked).view(1)
    return total, (stacked if per_tensor else None)


def multi_tensor_scale(noop_flag, tensor_lists, scale):
    src, dst = tensor_lists
    if not src:
        return
    if all(s.data_ptr() == d.data_ptr() for s, d in zip(src, dst)):
        torch._foreach_mul_(dst, scale)
        return
    for s, d in zip(src, dst):
        d.copy_(s.to(d.dtype) * scale if scale != 1.0 else s)
