"""Named contiguous scratch buffers (megatron/memory.py:24-132 surface, legacy)."""
import torch

_MEM_BUFFS = dict()


def allocate_mem_buff(name, numel, dtype, track_usage):
    assert name not in _MEM_BUFFS, f"memory buffer {name} already allocated."
    _MEM_BUFFS[name] = MemoryBuffer(name, numel, dtype, track_usage)
    return _MEM_BUFFS[name]


def get_mem_buff(name):
    return _MEM_BUFFS[name]


class MemoryBuffer:
    def __init__(self, name, numel, dtype, track_usage):
        self.name, self.numel, self.dtype = name, numel, dtype
        dev = torch.cuda.current_device() if torch.cuda.is_available() else "cpu"
        self.data = torch.empty(numel, dtype=dtype, device=dev, requires_grad=False)
        self._start = 0
        self.track_usage = track_usage
        self.in_use_value, self.total_value = 0.0, 0.0

    def reset(self):
        self._start = 0

    def is_in_use(self):
        return self._start > 0

    def numel_in_use(self):
        return self._start

    def add(self, tensor):
        assert tensor.dtype == self.dtype
        n = tensor.numel()
        new_start = self._start + n
        assert new_start <= self.numel, "Not enough memory left in the buffer"
        view = self.data[self._start:new_start].view(tensor.shape)
        view.copy_(tensor)
        self._start = new_start
        return view

    def get_data(self):
        if self.track_usage:
            self.in_use_value += float(self._start)
            self.total_value += float(self.numel)
        return self.data[:self._start]

    def print_average_usage(self):
        assert self.track_usage, "You need to enable track usage."
        print(" > usage of {} memory buffer: {:.2f} %".format(
            self.name, self.in_use_value * 100.0 / max(self.total_value, 1.0)), flush=True)


class RingMemBuffer:
    def __init__(self, name, num_buffers, numel, dtype, track_usage):
        self.num_buffers = num_buffers
        self.buffers = [allocate_mem_buff(f"{name} {i}", numel, dtype, track_usage) for i in range(num_buffers)]
        self._index = -1

    def get_next_buffer(self):
        self._index = (self._index + 1) % self.num_buffers
        buff = self.buffers[self._index]
        assert not buff.is_in_use(), "buffer is already in use."
        return buff
