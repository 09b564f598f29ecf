"""Environment / flag driven fault injection, used only by tests.

``--fault_inject kill:<rank>:<iteration>`` makes that rank raise SIGTERM at the start of
the iteration (exercises --exit_signal_handler checkpoint-and-exit);
``--fault_inject nan:<rank>:<iteration>`` poisons one gradient element (exercises the
skipped-iteration path)."""
import os
import signal

import torch.distributed as dist


def parse(spec):
    if not spec:
        return None
    kind, rank, it = spec.split(":")
    return kind, int(rank), int(it)


def maybe_inject(spec, iteration, models=None):
    f = parse(spec or os.environ.get("B200_FAULT_INJECT"))
    if f is None:
        return
    kind, rank, it = f
    me = dist.get_rank() if dist.is_initialized() else 0
    if me != rank or iteration != it:
        return
    if kind == "kill":
        os.kill(os.getpid(), signal.SIGTERM)
    elif kind == "nan" and models:
        for p in models[0].parameters():
            if hasattr(p, "main_grad"):
                p.main_grad.view(-1)[0] = float("nan")
                break
