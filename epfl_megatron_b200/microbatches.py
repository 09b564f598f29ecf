"""Number of micro-batches per step, constant or ramped up
(megatron/microbatches.py:59-144 behaviour)."""
from __future__ import annotations


def build_num_microbatches_calculator(args):
    if args.rampup_batch_size is None:
        calc = ConstantNumMicroBatches(args.global_batch_size, args.micro_batch_size,
                                       args.data_parallel_size)
        if args.rank == 0:
            print(f"setting number of micro-batches to constant {calc.get()}", flush=True)
        return calc
    assert len(args.rampup_batch_size) == 3, \
        "expected: --rampup_batch_size <start batch size> <batch size increment> <ramp-up samples>"
    start, incr, samples = (int(v) for v in args.rampup_batch_size)
    if args.rank == 0:
        print(f"will use batch size rampup starting from global batch size {start} to global batch "
              f"size {args.global_batch_size} with batch size increments {incr} over {samples} samples.",
              flush=True)
    return RampupBatchsizeNumMicroBatches(start, incr, samples, args.global_batch_size,
                                          args.micro_batch_size, args.data_parallel_size)


class NumMicroBatchesCalculator:
    def __init__(self):
        self.num_micro_batches = None
        self.current_global_batch_size = None

    def get(self):
        return self.num_micro_batches

    def get_current_global_batch_size(self):
        return self.current_global_batch_size

    def update(self, consumed_samples, consistency_check):
        pass


class ConstantNumMicroBatches(NumMicroBatchesCalculator):
    def __init__(self, global_batch_size, micro_batch_size, data_parallel_size):
        super().__init__()
        per_step = micro_batch_size * data_parallel_size
        assert global_batch_size % per_step == 0, \
            f"global batch size ({global_batch_size}) is not divisible by micro batch size " \
            f"({micro_batch_size}) times data parallel size ({data_parallel_size})"
        self.num_micro_batches = global_batch_size // per_step
        assert self.num_micro_batches >= 1
        self.current_global_batch_size = global_batch_size


class RampupBatchsizeNumMicroBatches(NumMicroBatchesCalculator):
    def __init__(self, start_batch_size, batch_size_increment, ramup_samples, global_batch_size,
                 micro_batch_size, data_parallel_size):
        super().__init__()
        self.micro_batch_size = micro_batch_size
        self.data_parallel_size = data_parallel_size
        self.per_step = micro_batch_size * data_parallel_size
        assert self.per_step > 0 and start_batch_size > 0 and global_batch_size > 0
        self.start_batch_size = start_batch_size
        self.global_batch_size = global_batch_size
        diff = global_batch_size - start_batch_size
        assert diff >= 0 and batch_size_increment > 0
        self.batch_size_increment = batch_size_increment
        assert diff % batch_size_increment == 0, \
            "expected global batch size interval to be divisible by the batch size increment"
        n_incr = diff // batch_size_increment
        self.ramup_samples = ramup_samples
        assert ramup_samples >= 0
        self.rampup_samples_per_increment = ramup_samples / max(n_incr, 1)
        self.update(0, False)

    def update(self, consumed_samples, consistency_check):
        if consumed_samples > self.ramup_samples:
            self.current_global_batch_size = self.global_batch_size
        else:
            steps = int(consumed_samples / self.rampup_samples_per_increment)
            self.current_global_batch_size = self.start_batch_size + steps * self.batch_size_increment
            assert self.current_global_batch_size <= self.global_batch_size
        if consistency_check:
            assert self.current_global_batch_size % self.per_step == 0, \
                f"current global batch size ({self.current_global_batch_size}) is not divisible by " \
                f"micro-batch-size ({self.micro_batch_size}) times data parallel size ({self.data_parallel_size})"
        self.num_micro_batches = self.current_global_batch_size // self.per_step
