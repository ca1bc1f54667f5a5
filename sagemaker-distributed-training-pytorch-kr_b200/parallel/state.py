"""Model-parallel topology: which ranks form the tensor-, pipeline-, data-, model-parallel and
embedding groups, plus the per-process state (group handles, ranks, virtual-pipeline rank).

Plays the role of megatron.core.parallel_state (used in-tree at gpt_dataset.py:463-467 and by
every tensor-parallel layer).  Rank layout rule (same as the reference stack so checkpoints and logs line
up): tensor-parallel ranks are consecutive, data-parallel ranks stride by TP inside a pipeline stage,
pipeline ranks stride by world/PP.  ``compute_layout`` is a pure function so it can be table-tested.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, List, Optional

import torch
import torch.distributed as dist


@dataclass
class Layout:
    world: int
    tp: int
    pp: int
    dp: int
    tp_groups: List[List[int]] = field(default_factory=list)
    pp_groups: List[List[int]] = field(default_factory=list)
    dp_groups: List[List[int]] = field(default_factory=list)
    mp_groups: List[List[int]] = field(default_factory=list)
    embedding_groups: List[List[int]] = field(default_factory=list)
    position_embedding_groups: List[List[int]] = field(default_factory=list)


def compute_layout(world: int, tp: int = 1, pp: int = 1, split_rank: Optional[int] = None) -> Layout:
    if world % (tp * pp) != 0:
        raise ValueError(f"world size {world} not divisible by tp {tp} x pp {pp}")
    dp = world // (tp * pp)
    lay = Layout(world, tp, pp, dp)
    per_stage = world // pp
    # data-parallel: same tp index, same stage
    for stage in range(pp):
        lo = stage * per_stage
        for t in range(tp):
            lay.dp_groups.append(list(range(lo + t, lo + per_stage, tp)))
    # model-parallel: everything that shares a dp index
    for d in range(dp):
        grp = []
        for stage in range(pp):
            lo = stage * per_stage + d * tp
            grp.extend(range(lo, lo + tp))
        lay.mp_groups.append(sorted(grp))
    # tensor-parallel: consecutive
    for i in range(world // tp):
        lay.tp_groups.append(list(range(i * tp, (i + 1) * tp)))
    # pipeline: stride world/pp ; embedding = first + last stage (tied embeddings), position = first
    for i in range(per_stage):
        ranks = list(range(i, world, per_stage))
        lay.pp_groups.append(ranks)
        if len(ranks) > 1:
            emb, pos = [ranks[0], ranks[-1]], [ranks[0]]
            if split_rank is not None:
                if ranks[split_rank] not in emb:
                    emb = [ranks[0], ranks[split_rank], ranks[-1]]
                if ranks[split_rank] not in pos:
                    pos = [ranks[0], ranks[split_rank]]
        else:
            emb, pos = ranks, ranks
        lay.embedding_groups.append(emb)
        lay.position_embedding_groups.append(pos)
    return lay


class _State:
    def __init__(self):
        self.reset()

    def reset(self):
        self.initialized = False
        self.layout: Optional[Layout] = None
        self.rank = 0
        self.groups: Dict[str, object] = {}
        self.group_ranks: Dict[str, List[int]] = {}
        self.virtual_pp_rank: Optional[int] = None
        self.virtual_pp_size: Optional[int] = None
        self.split_rank: Optional[int] = None
        self.gloo_dp_group = None


_S = _State()


def initialize_model_parallel(tensor_model_parallel_size: int = 1, pipeline_model_parallel_size: int = 1,
                              virtual_pipeline_model_parallel_size: Optional[int] = None,
                              pipeline_model_parallel_split_rank: Optional[int] = None, verbose: bool = False) -> Layout:
    """Creates every process group (all ranks must call this; new_group is collective)."""
    world = dist.get_world_size() if dist.is_initialized() else 1
    rank = dist.get_rank() if dist.is_initialized() else 0
    lay = compute_layout(world, tensor_model_parallel_size, pipeline_model_parallel_size, pipeline_model_parallel_split_rank)
    _S.reset()
    _S.layout, _S.rank = lay, rank
    _S.virtual_pp_size = virtual_pipeline_model_parallel_size
    _S.virtual_pp_rank = 0 if virtual_pipeline_model_parallel_size else None
    _S.split_rank = pipeline_model_parallel_split_rank
    if verbose and rank == 0:
        print(f"> initialized tensor model parallel with size {lay.tp}", flush=True)
        print(f"> initialized pipeline model parallel with size {lay.pp}", flush=True)

    def build(name: str, rank_lists: List[List[int]]):
        mine, mine_ranks = None, None
        for ranks in rank_lists:
            g = dist.new_group(ranks) if (dist.is_initialized() and world > 1) else None
            if rank in ranks:
                mine, mine_ranks = g, ranks
        _S.groups[name], _S.group_ranks[name] = mine, mine_ranks

    build("dp", lay.dp_groups)
    build("mp", lay.mp_groups)
    build("tp", lay.tp_groups)
    build("pp", lay.pp_groups)
    build("embedding", lay.embedding_groups)
    build("position_embedding", lay.position_embedding_groups)
    _S.initialized = True
    return lay


def destroy_model_parallel():
    _S.reset()


def model_parallel_is_initialized() -> bool:
    return _S.initialized


def _g(name):
    if not _S.initialized:
        return None
    return _S.groups.get(name)


def _ranks(name) -> List[int]:
    if not _S.initialized or _S.group_ranks.get(name) is None:
        return [0]
    return _S.group_ranks[name]


def get_tensor_model_parallel_group(): return _g("tp")
def get_pipeline_model_parallel_group(): return _g("pp")
def get_data_parallel_group(): return _g("dp")
def get_model_parallel_group(): return _g("mp")
def get_embedding_group(): return _g("embedding")
def get_position_embedding_group(): return _g("position_embedding")

def get_tensor_model_parallel_ranks(): return _ranks("tp")
def get_pipeline_model_parallel_ranks(): return _ranks("pp")
def get_data_parallel_ranks(): return _ranks("dp")

def get_tensor_model_parallel_world_size() -> int: return len(_ranks("tp"))
def get_pipeline_model_parallel_world_size() -> int: return len(_ranks("pp"))
def get_data_parallel_world_size() -> int: return len(_ranks("dp"))

def get_tensor_model_parallel_rank() -> int: return _ranks("tp").index(_S.rank) if _S.initialized else 0
def get_pipeline_model_parallel_rank() -> int: return _ranks("pp").index(_S.rank) if _S.initialized else 0
def get_data_parallel_rank() -> int: return _ranks("dp").index(_S.rank) if _S.initialized else 0

def get_tensor_model_parallel_src_rank() -> int: return _ranks("tp")[0]
def get_data_parallel_src_rank() -> int: return _ranks("dp")[0]
def get_pipeline_model_parallel_first_rank() -> int: return _ranks("pp")[0]
def get_pipeline_model_parallel_last_rank() -> int: return _ranks("pp")[-1]


def get_pipeline_model_parallel_next_rank() -> int:
    r = _ranks("pp")
    return r[(get_pipeline_model_parallel_rank() + 1) % len(r)]


def get_pipeline_model_parallel_prev_rank() -> int:
    r = _ranks("pp")
    return r[(get_pipeline_model_parallel_rank() - 1) % len(r)]


def get_virtual_pipeline_model_parallel_rank(): return _S.virtual_pp_rank
def get_virtual_pipeline_model_parallel_world_size(): return _S.virtual_pp_size
def set_virtual_pipeline_model_parallel_rank(r): _S.virtual_pp_rank = r


def is_pipeline_first_stage(ignore_virtual: bool = False) -> bool:
    if not ignore_virtual and _S.virtual_pp_size is not None and _S.virtual_pp_rank != 0:
        return False
    return get_pipeline_model_parallel_rank() == 0


def is_pipeline_last_stage(ignore_virtual: bool = False) -> bool:
    if not ignore_virtual and _S.virtual_pp_size is not None and _S.virtual_pp_rank != _S.virtual_pp_size - 1:
        return False
    return get_pipeline_model_parallel_rank() == get_pipeline_model_parallel_world_size() - 1


def is_rank_in_embedding_group(ignore_virtual: bool = False) -> bool:
    ranks = _ranks("embedding")
    if _S.rank not in ranks:
        return False
    if ignore_virtual:
        return True
    if _S.rank == ranks[0]:
        return is_pipeline_first_stage()
    if _S.rank == ranks[-1]:
        return is_pipeline_last_stage()
    return True


def is_rank_in_position_embedding_group() -> bool:
    return _S.rank in _ranks("position_embedding")


# ------------------------------------------------------------------------------------------
# remaining pieces of megatron.core.parallel_state
# ------------------------------------------------------------------------------------------
def get_data_parallel_group_gloo():
    """CPU (gloo) twin of the data-parallel group, used for host-side exchanges of the distributed optimizer.  Created on first
    use; like every `new_group` this is collective over the world: all ranks must make the first call together."""
    if not _S.initialized:
        return None
    if "dp_gloo" not in _S.groups:
        mine = None
        if dist.is_initialized() and dist.get_world_size() > 1:
            for ranks in _S.layout.dp_groups:
                g = dist.new_group(ranks, backend="gloo")
                if _S.rank in ranks:
                    mine = g
        _S.groups["dp_gloo"], _S.group_ranks["dp_gloo"] = mine, _S.group_ranks["dp"]
    return _S.groups["dp_gloo"]


def get_pipeline_model_parallel_split_rank() -> Optional[int]:
    return _S.split_rank


def is_pipeline_stage_before_split(rank: Optional[int] = None) -> bool:
    """encoder side of an encoder-decoder pipeline (`--pipeline-model-parallel-split-rank`); everything counts as "before" without a split"""
    if get_pipeline_model_parallel_world_size() == 1 or _S.split_rank is None:
        return True
    r = get_pipeline_model_parallel_rank() if rank is None else rank
    return r < _S.split_rank


def is_pipeline_stage_after_split(rank: Optional[int] = None) -> bool:
    if get_pipeline_model_parallel_world_size() == 1 or _S.split_rank is None:
        return True
    r = get_pipeline_model_parallel_rank() if rank is None else rank
    return r >= _S.split_rank


def is_pipeline_stage_at_split() -> bool:
    r = get_pipeline_model_parallel_rank()
    return is_pipeline_stage_before_split(r) and is_pipeline_stage_after_split(r + 1)


class GlobalMemoryBuffer:
    """Named, growing scratch tensors reused across calls (megatron.core.utils.GlobalMemoryBuffer): `get_tensor(shape, dtype, name)`
    returns a view of a buffer that is reallocated only when a larger request arrives."""

    def __init__(self):
        self.buffer = {}

    def get_tensor(self, tensor_shape, dtype, name, device=None):
        import math
        import torch
        need = math.prod(tensor_shape)
        key = (name, dtype)
        buf = self.buffer.get(key)
        if buf is None or buf.numel() < need or (device is not None and buf.device != torch.device(device)):
            dev = device if device is not None else (torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else "cpu")
            buf = torch.empty(need, dtype=dtype, device=dev, requires_grad=False)
            self.buffer[key] = buf
        return buf[:need].view(*tensor_shape)


_GLOBAL_MEMORY_BUFFER = GlobalMemoryBuffer()


def get_global_memory_buffer() -> GlobalMemoryBuffer:
    return _GLOBAL_MEMORY_BUFFER
