"""DeepSpeed ZeRO-3 checkpoint layout (produced by the reference at train.py:246 because
`stage3_gather_16bit_weights_on_model_save` is false, log 4_...ipynb:2311):

    <dir>/latest                                          -> "global_step{N}"
    <dir>/global_step{N}/zero_pp_rank_{r}_mp_rank_00_model_states.pt         (module placeholders, param_shapes, ds_config, ...)
    <dir>/global_step{N}/bf16_zero_pp_rank_{r}_mp_rank_00_optim_states.pt    (zero_stage 3, partition_count, fp32_flat_groups, Adam state)

Our partition is per gather unit (tile-aligned, padded) rather than DeepSpeed's single flat buffer; the per-unit
offsets and shapes are stored next to the shards so ``zero_to_fp32`` can consolidate any world size's files into one
fp32 state_dict (the role of DeepSpeed's zero_to_fp32.py).
"""
from __future__ import annotations

import os
from typing import Dict, List, Optional, Tuple

import torch
import torch.distributed as dist

DS_VERSION = "b200dist-zero3-1"


def _names(step: int, rank: int, bf16: bool = True) -> Tuple[str, str]:
    return (f"zero_pp_rank_{rank}_mp_rank_00_model_states.pt",
            f"{'bf16_' if bf16 else ''}zero_pp_rank_{rank}_mp_rank_00_optim_states.pt")


def save_zero3_checkpoint(engine, out_dir: str, step: int, ds_config: Dict, scheduler=None, persistent=None):
    tag = f"global_step{step}"
    d = os.path.join(out_dir, tag)
    os.makedirs(d, exist_ok=True)
    st = engine.shard_state()
    bf16 = engine.dtype == torch.bfloat16
    mfile, ofile = _names(step, engine.rank, bf16)
    model_states = {
        "module": {n: p.detach().cpu() for p, n in (persistent or [])},       # replicated params; partitioned ones are placeholders
        "param_shapes": st["param_shapes"], "param_offsets": st["param_offsets"], "unit_names": st["unit_names"],
        "unit_numel": st["unit_numel"], "shared_params": {}, "ds_config": ds_config, "ds_version": DS_VERSION,
        "global_steps": step, "dp_world_size": engine.world, "mp_world_size": 1,
        "lr_scheduler": scheduler.state_dict() if scheduler is not None else None,
    }
    optim_states = {"optimizer_state_dict": {
        "zero_stage": 3, "partition_count": engine.world, "fp32_flat_groups": st["fp32_flat_groups"],
        "optimizer_state_dict": {"state": {i: {"exp_avg": st["exp_avg"][i], "exp_avg_sq": st["exp_avg_sq"][i], "step": st["step"]}
                                           for i in range(len(st["fp32_flat_groups"]))}},
        "persistent_state": {k: st.get(k) for k in ("persist_fp32", "persist_m", "persist_v")},
        "rank": engine.rank}, "ds_config": ds_config, "ds_version": DS_VERSION}
    torch.save(model_states, os.path.join(d, mfile))
    torch.save(optim_states, os.path.join(d, ofile))
    if engine.rank == 0:
        with open(os.path.join(out_dir, "latest"), "w") as f:
            f.write(tag)
        # DeepSpeed leaves its consolidation script next to the shards; so do we (tools/zero_to_fp32.py, torch only)
        src = os.path.join(os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))), "tools", "zero_to_fp32.py")
        if os.path.isfile(src):
            import shutil
            shutil.copyfile(src, os.path.join(out_dir, "zero_to_fp32.py"))
    if dist.is_initialized():
        dist.barrier()


def load_zero3_checkpoint(engine, ckpt_dir: str, tag: Optional[str] = None, scheduler=None) -> int:
    if tag is None:
        tag = open(os.path.join(ckpt_dir, "latest")).read().strip()
    d = os.path.join(ckpt_dir, tag)
    step = int(tag.replace("global_step", ""))
    bf16 = engine.dtype == torch.bfloat16
    mfile, ofile = _names(step, engine.rank, bf16)
    ms = torch.load(os.path.join(d, mfile), map_location="cpu", weights_only=False)
    os_ = torch.load(os.path.join(d, ofile), map_location="cpu", weights_only=False)["optimizer_state_dict"]
    st = os_["optimizer_state_dict"]["state"]
    engine.load_shard_state({"partition_count": os_["partition_count"], "step": st[0]["step"] if st else 0,
                             "fp32_flat_groups": os_["fp32_flat_groups"], "exp_avg": [st[i]["exp_avg"] for i in range(len(st))],
                             "exp_avg_sq": [st[i]["exp_avg_sq"] for i in range(len(st))], **(os_.get("persistent_state") or {})})
    if scheduler is not None and ms.get("lr_scheduler"):
        scheduler.load_state_dict(ms["lr_scheduler"])
    return ms.get("global_steps", step)


def zero_to_fp32(ckpt_dir: str, tag: Optional[str] = None) -> Dict[str, torch.Tensor]:
    """Consolidate all ranks' shards into one fp32 state_dict keyed `<unit>.<param>` (+ the replicated params)."""
    if tag is None:
        tag = open(os.path.join(ckpt_dir, "latest")).read().strip()
    d = os.path.join(ckpt_dir, tag)
    files = sorted(f for f in os.listdir(d) if f.endswith("_optim_states.pt"))
    world = len(files)
    shards: List[Optional[Dict]] = [None] * world
    for f in files:
        o = torch.load(os.path.join(d, f), map_location="cpu", weights_only=False)["optimizer_state_dict"]
        shards[o["rank"]] = o
    assert all(s is not None for s in shards) and shards[0]["partition_count"] == world, "missing rank shards"
    mfile = sorted(f for f in os.listdir(d) if f.endswith("_model_states.pt"))[0]
    ms = torch.load(os.path.join(d, mfile), map_location="cpu", weights_only=False)
    out: Dict[str, torch.Tensor] = {}
    for u, uname in enumerate(ms["unit_names"]):
        flat = torch.cat([shards[r]["fp32_flat_groups"][u] for r in range(world)])       # rank order == slice order
        for pname, shape in ms["param_shapes"][u].items():
            off = ms["param_offsets"][u][pname]
            n = 1
            for s_ in shape:
                n *= s_
            out[f"{uname}.{pname}"] = flat[off:off + n].view(shape).clone()
    for n, t in ms["module"].items():
        out[n] = t.float()
    return out
