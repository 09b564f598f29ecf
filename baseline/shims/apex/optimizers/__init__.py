import torch


class FusedAdam(torch.optim.AdamW):
    """apex.optimizers.FusedAdam(adam_w_mode=True) -> torch.optim.AdamW(fused=True)."""

    def __init__(self, params, lr=1e-3, bias_correction=True, betas=(0.9, 0.999), eps=1e-8,
                 adam_w_mode=True, weight_decay=0.0, amsgrad=False, set_grad_none=True, **kw):
        assert adam_w_mode and not amsgrad
        super().__init__(params, lr=lr, betas=betas, eps=eps, weight_decay=weight_decay,
                         fused=torch.cuda.is_available())


class FusedSGD(torch.optim.SGD):
    def __init__(self, params, lr=1e-3, momentum=0.0, dampening=0.0, weight_decay=0.0, nesterov=False, **kw):
        super().__init__(params, lr=lr, momentum=momentum, dampening=dampening, weight_decay=weight_decay,
                         nesterov=nesterov)
