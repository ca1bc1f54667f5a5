"""Number-of-micro-batches calculators (megatron/microbatches.py, part of SURVEY E8): constant, or ramped up with
`--rampup-batch-size <start> <increment> <ramp-up samples>` - the global batch grows from <start> to --global-batch-size in
steps of <increment>, spread evenly over the first <ramp-up samples> samples (sample-based training only).

    build_num_microbatches_calculator(args); get_num_microbatches(); get_current_global_batch_size();
    update_num_microbatches(consumed_samples, consistency_check=True)
"""
from __future__ import annotations

from typing import Optional


class ConstantNumMicroBatches:
    def __init__(self, global_batch_size: int, micro_batch_size: int, data_parallel_size: int):
        per = micro_batch_size * data_parallel_size
        assert global_batch_size % per == 0, "global batch size ({}) is not divisible by micro batch size ({}) times data parallel size ({})".format(
            global_batch_size, micro_batch_size, data_parallel_size)
        self.num_micro_batches = global_batch_size // per
        assert self.num_micro_batches >= 1
        self.current_global_batch_size = global_batch_size

    def get(self) -> int:
        return self.num_micro_batches

    def get_current_global_batch_size(self) -> int:
        return self.current_global_batch_size

    def update(self, consumed_samples: int, consistency_check: bool):
        pass


class RampupBatchsizeNumMicroBatches(ConstantNumMicroBatches):
    def __init__(self, start_batch_size: int, batch_size_increment: int, ramup_samples: int, global_batch_size: int, micro_batch_size: int,
                 data_parallel_size: int):
        self.micro_batch_size, self.data_parallel_size = micro_batch_size, data_parallel_size
        self.per = micro_batch_size * data_parallel_size
        assert self.per > 0 and start_batch_size > 0 and batch_size_increment > 0 and ramup_samples >= 0 and global_batch_size > 0
        self.start_batch_size, self.batch_size_increment = start_batch_size, batch_size_increment
        self.ramup_samples, self.global_batch_size = ramup_samples, global_batch_size
        diff = global_batch_size - start_batch_size
        assert diff >= 0, "expected global batch size to be greater than or equal to start batch size, got {} and {}".format(global_batch_size, start_batch_size)
        assert diff % batch_size_increment == 0, "expected global batch size interval ({}) to be divisible by global batch size increment ({})".format(
            diff, batch_size_increment)
        num_increments = diff // batch_size_increment
        self.rampup_samples_per_increment = self.ramup_samples / num_increments if num_increments else float("inf")
        self.update(0, False)

    def update(self, consumed_samples: int, consistency_check: bool):
        if consumed_samples > self.ramup_samples:
            self.current_global_batch_size = self.global_batch_size
        else:
            steps = int(consumed_samples / self.rampup_samples_per_increment)
            self.current_global_batch_size = self.start_batch_size + steps * self.batch_size_increment
            assert self.current_global_batch_size <= self.global_batch_size
        if consistency_check:
            assert self.current_global_batch_size % self.per == 0, \
                "current global batch size ({}) is not divisible by micro-batch-size ({}) times data parallel size ({})".format(
                    self.current_global_batch_size, self.micro_batch_size, self.data_parallel_size)
        self.num_micro_batches = self.current_global_batch_size // self.per


_CALC: Optional[ConstantNumMicroBatches] = None


def build_num_microbatches_calculator(args, verbose: bool = True):
    global _CALC
    if args.rampup_batch_size is None:
        _CALC = ConstantNumMicroBatches(args.global_batch_size, args.micro_batch_size, args.data_parallel_size)
        if verbose and args.rank == 0:
            print("setting number of micro-batches to constant {}".format(_CALC.get()), flush=True)
    else:
        assert len(args.rampup_batch_size) == 3, "expected the following format: --rampup-batch-size <start batch size> <batch size incerement> <ramp-up samples>"
        start, inc, ramp = (int(x) for x in args.rampup_batch_size)
        if verbose and args.rank == 0:
            print("will use batch size rampup starting from global batch size {} to global batch size {} with batch size increments {} over {} samples.".format(
                start, args.global_batch_size, inc, ramp), flush=True)
        _CALC = RampupBatchsizeNumMicroBatches(start, inc, ramp, args.global_batch_size, args.micro_batch_size, args.data_parallel_size)
    return _CALC


def get_num_microbatches() -> int:
    return _CALC.get()


def get_current_global_batch_size() -> int:
    return _CALC.get_current_global_batch_size()


def update_num_microbatches(consumed_samples: int, consistency_check: bool = True):
    _CALC.update(consumed_samples, consistency_check)
