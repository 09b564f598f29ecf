"""Named timers (megatron/timers.py:56-307 surface: ``timers(name, log_level).start()``,
``.stop()``, ``.log(names)``, ``.write(...)``).

Device time is measured with CUDA events recorded on the current stream, so starting and
stopping a timer does not drain the GPU (the reference calls torch.cuda.synchronize() on
every start/stop). Elapsed values are resolved lazily when read; the cross-rank
min/max report all-gathers one float per timer.
"""
from __future__ import annotations

import time

import torch
import torch.distributed as dist


class _Base:
    def __init__(self, name):
        self.name = name

    def start(self, barrier=False):
        pass

    def stop(self, barrier=False):
        pass

    def reset(self):
        pass

    def elapsed(self, reset=True, barrier=False):
        return 0.0


class DummyTimer(_Base):
    def __init__(self):
        super().__init__("dummy")

    def elapsed(self, reset=True, barrier=False):
        raise Exception("dummy timer should not be used to calculate elapsed time")


# --nvtx_ranges: every timer interval is also an NVTX range (visible in nsys / ncu --nvtx); the
# reference has no profiler integration at all.
_NVTX = False


def set_nvtx_ranges(on: bool):
    global _NVTX
    _NVTX = bool(on) and torch.cuda.is_available()


class Timer(_Base):
    def __init__(self, name):
        super().__init__(name)
        self._elapsed = 0.0
        self._started = False
        self._pending = []      # (start_event, stop_event) pairs not yet resolved
        self._start_evt = None
        self._t0 = 0.0
        self._barrier_group = None
        self._cuda = torch.cuda.is_available()

    def set_barrier_group(self, group):
        self._barrier_group = group

    def start(self, barrier=False):
        assert not self._started, "timer has already been started"
        if barrier and dist.is_initialized():
            dist.barrier(group=self._barrier_group)
        if self._cuda:
            self._start_evt = torch.cuda.Event(enable_timing=True)
            self._start_evt.record()
        else:
            self._t0 = time.time()
        if _NVTX:
            torch.cuda.nvtx.range_push(self.name)
        self._started = True

    def stop(self, barrier=False):
        assert self._started, "timer is not started"
        if _NVTX:
            torch.cuda.nvtx.range_pop()
        if barrier and dist.is_initialized():
            dist.barrier(group=self._barrier_group)
        if self._cuda:
            e = torch.cuda.Event(enable_timing=True)
            e.record()
            self._pending.append((self._start_evt, e))
        else:
            self._elapsed += time.time() - self._t0
        self._started = False

    def _resolve(self):
        for s, e in self._pending:
            e.synchronize()
            self._elapsed += s.elapsed_time(e) / 1000.0
        self._pending = []

    def reset(self):
        self._elapsed = 0.0
        self._pending = []
        self._started = False

    def elapsed(self, reset=True, barrier=False):
        was = self._started
        if was:
            self.stop(barrier=barrier)
        self._resolve()
        val = self._elapsed
        if reset:
            self.reset()
        if was:
            self.start(barrier=barrier)
        return val


class Timers:
    def __init__(self, log_level, log_option):
        self._log_level = log_level
        self._log_option = log_option
        self._timers = {}
        self._log_levels = {}
        self._dummy = DummyTimer()
        self._max_log_level = 2

    def __call__(self, name, log_level=None):
        if name in self._timers:
            if log_level is not None:
                assert log_level == self._log_levels[name], \
                    f"input log level {log_level} does not match already existing log level " \
                    f"{self._log_levels[name]} for {name} timer"
            return self._timers[name]
        if log_level is None:
            log_level = self._max_log_level
        assert log_level <= self._max_log_level
        if log_level > self._log_level:
            return self._dummy
        self._timers[name] = Timer(name)
        self._log_levels[name] = log_level
        return self._timers[name]

    def _gather(self, names, reset, barrier):
        if barrier and dist.is_initialized():
            dist.barrier()
        world = dist.get_world_size() if dist.is_initialized() else 1
        rank = dist.get_rank() if dist.is_initialized() else 0
        use_cuda = torch.cuda.is_available() and (not dist.is_initialized() or dist.get_backend() != "gloo")
        dev = torch.cuda.current_device() if use_cuda else "cpu"
        local = torch.zeros(len(names), dtype=torch.float, device=dev)
        for i, n in enumerate(names):
            if n in self._timers:
                local[i] = self._timers[n].elapsed(reset=reset)
        if world == 1:
            return local.view(1, -1).cpu()
        table = [torch.zeros_like(local) for _ in range(world)]
        dist.all_gather(table, local)
        return torch.stack(table).cpu()

    def _report(self, names, reset, barrier, normalizer):
        assert normalizer > 0.0
        table = self._gather(names, reset, barrier) / normalizer
        opt = self._log_option
        lines = []
        if opt in ("max", "minmax"):
            head = "(min, max) time across ranks (ms):" if opt == "minmax" else "max time across ranks (ms):"
            rows = []
            for i, n in enumerate(names):
                col = table[:, i]
                col = col[col > 0.0]
                if col.numel() == 0:
                    continue
                if opt == "minmax":
                    rows.append("    {}: ({:.2f}, {:.2f})".format(
                        (n + " ").ljust(48, "."), float(col.min()) * 1e3, float(col.max()) * 1e3))
                else:
                    rows.append("    {}: {:.2f}".format((n + " ").ljust(48, "."), float(col.max()) * 1e3))
            if rows:
                lines = [head] + rows
        elif opt == "all":
            lines = ["times across ranks (ms):"]
            for i, n in enumerate(names):
                got = False
                for r in range(table.shape[0]):
                    if table[r, i] > 0:
                        if not got:
                            lines.append(f"  {n}:")
                            got = True
                        lines.append("     rank {:2d}: {:.2f}".format(r, float(table[r, i]) * 1e3))
        else:
            raise Exception(f"unknown timing log option {opt}")
        return "\n".join(lines) if lines else None

    def log(self, names, rank=None, normalizer=1.0, reset=True, barrier=False):
        out = self._report(names, reset, barrier, normalizer)
        if rank is None:
            rank = (dist.get_world_size() - 1) if dist.is_initialized() else 0
        me = dist.get_rank() if dist.is_initialized() else 0
        if me == rank and out is not None:
            print(out, flush=True)

    def write(self, names, writer, iteration, normalizer=1.0, reset=False, barrier=False):
        assert normalizer > 0.0
        table = self._gather(names, reset, barrier)
        if writer is not None:
            for i, n in enumerate(names):
                writer.add_scalar(n + "-time", float(table[:, i].max()) / normalizer, iteration)
