"""Pipeline-parallel engine: schedule generation + dataflow executor over NCCL (or gloo) p2p.

Counterpart of megatron.core.pipeline_parallel.{schedules,p2p_communication} (C18):
``get_forward_backward_func()`` returns one of forward_backward_no_pipelining / 1F1B / interleaved 1F1B.

Design:
  * The per-rank compute order comes from the C++ generator (csrc/host/pipeline_schedule.cpp).
  * Communication is derived from that order: a forward step on a non-first virtual stage consumes one
    activation from the previous virtual stage, and produces one for the next; likewise for gradients.
    Each DIRECTED neighbour link (rank -> next for activations, rank -> prev for gradients, including the
    wrap-around links the interleaved schedule uses) gets its own 2-rank process group, so every
    communicator is used in one direction only; with FIFO order per link there is no send/recv cycle and
    no need for the reference's fused send_forward_recv_backward calls.
  * Receives for the next step are posted before the current step computes, sends are asynchronous:
    p2p runs on the communicator's own stream and overlaps with compute (the `--overlap-p2p-communication`
    behaviour is the default here).
  * Tensor shape on the wire: [s (/t under sequence parallelism), b, h] in the pipeline dtype.
"""
from __future__ import annotations

import contextlib
import os
import time
import ctypes
from typing import Callable, Dict, Iterator, List, Optional, Tuple, Union

import torch
import torch.distributed as dist

from ..utils import hostlib
from . import state as mpu

FWD, BWD = 0, 1


def build_schedule(P: int, rank: int, V: int, M: int, forward_only: bool = False) -> List[Tuple[int, int, int]]:
    """[(op, chunk, microbatch)] for this rank (C++ generator)."""
    cap = 2 * M * V + 8
    buf = (ctypes.c_int32 * (3 * cap))()
    n = hostlib.lib().b2h_pp_schedule(ctypes.c_int32(P), ctypes.c_int32(rank), ctypes.c_int32(V), ctypes.c_int32(M),
                                      ctypes.c_int32(int(forward_only)), buf, ctypes.c_int32(cap))
    if n == -1:
        raise ValueError(f"invalid pipeline schedule request P={P} rank={rank} V={V} M={M} "
                         "(interleaving needs num_microbatches % pipeline size == 0)")
    assert n >= 0
    return [(buf[3 * i], buf[3 * i + 1], buf[3 * i + 2]) for i in range(n)]


def build_schedule_py(P: int, rank: int, V: int, M: int, forward_only: bool = False):
    """Pure-Python twin of the C++ generator (test oracle)."""
    ops = []
    if V == 1:
        warm = min(P - rank - 1, M) if P > 1 else 0
        f = b = 0
        for _ in range(warm):
            ops.append((FWD, 0, f)); f += 1
        for _ in range(M - warm):
            ops.append((FWD, 0, f)); f += 1
            if not forward_only:
                ops.append((BWD, 0, b)); b += 1
        while not forward_only and b < M:
            ops.append((BWD, 0, b)); b += 1
        return ops
    assert M % P == 0
    total = M * V
    chunk = lambda k, fw: (k % (P * V)) // P if fw else V - 1 - (k % (P * V)) // P
    mb = lambda k: (k // (P * V)) * P + k % P
    warm = total if (forward_only or M == P) else min((P - rank - 1) * 2 + (V - 1) * P, total)
    f = b = 0
    for _ in range(warm):
        ops.append((FWD, chunk(f, True), mb(f))); f += 1
    for _ in range(total - warm):
        ops.append((FWD, chunk(f, True), mb(f))); f += 1
        ops.append((BWD, chunk(b, False), mb(b))); b += 1
    while not forward_only and b < total:
        ops.append((BWD, chunk(b, False), mb(b))); b += 1
    return ops


def virtual_stage(rank: int, chunk: int, P: int) -> int:
    return chunk * P + rank


def simulate(P: int, V: int, M: int) -> Dict[str, object]:
    """Dependency simulation of all ranks' schedules: proves deadlock-freedom under the FIFO-per-link
    dataflow rule and reports the peak number of in-flight forward activations per rank."""
    scheds = [build_schedule_py(P, r, V, M) for r in range(P)]
    pos = [0] * P
    done_f, done_b = set(), set()
    inflight, peak = [0] * P, [0] * P
    last = P * V - 1
    progressed = True
    steps = 0
    while progressed:
        progressed = False
        for r in range(P):
            while pos[r] < len(scheds[r]):
                op, c, m = scheds[r][pos[r]]
                vs = virtual_stage(r, c, P)
                if op == FWD:
                    ready = vs == 0 or (vs - 1, m) in done_f
                else:
                    ready = (vs, m) in done_f and (vs == last or (vs + 1, m) in done_b)
                if not ready:
                    break
                if op == FWD:
                    done_f.add((vs, m)); inflight[r] += 1; peak[r] = max(peak[r], inflight[r])
                else:
                    done_b.add((vs, m)); inflight[r] -= 1
                pos[r] += 1
                progressed = True
                steps += 1
    ok = all(pos[r] == len(scheds[r]) for r in range(P))
    return {"ok": ok, "peak_inflight": peak, "steps": steps}


# ------------------------------------------------------------------------------------------
# batched p2p plan: ONE fused send+recv group per schedule step on one communicator per pipeline group
# ------------------------------------------------------------------------------------------
def outgoing_messages(P: int, V: int, M: int, src: int, forward_only: bool = False, sched=None):
    """Per schedule step of rank `src`: None, or (dst rank, key at dst) for the tensor that step produces for a neighbour.
    key = (op, chunk, microbatch) of the step that CONSUMES it on the destination."""
    last = P * V - 1
    out = []
    for op, c, m in (sched if sched is not None else build_schedule_py(P, src, V, M, forward_only)):
        vs = virtual_stage(src, c, P)
        if op == FWD and vs < last:
            out.append(((src + 1) % P, (FWD, (vs + 1) // P, m)))
        elif op == BWD and vs > 0:
            out.append(((src - 1) % P, (BWD, (vs - 1) // P, m)))
        else:
            out.append(None)
    return out


def group_boundaries(sched, V: int):
    """Steps after which a communication group is issued.  Plain 1F1B: after every step.  Interleaved: after every step
    of the warm-up and cool-down phases, and after each (forward, backward) PAIR of the steady state - the sends of both
    steps and the receives for the next pair travel in one group, like the reference's
    send_forward_backward_recv_forward_backward (schedules.py, interleaved 1F1B).  A group after the forward of a pair
    would couple that forward's send to the backward's receive and can deadlock on an in-order transport."""
    n = len(sched)
    lead = next((i for i, st in enumerate(sched) if st[0] == BWD), n)      # leading forwards
    total_f = sum(1 for st in sched if st[0] == FWD)
    if V == 1 or lead == total_f:           # plain 1F1B, forward-only, or "all warm-up" (num_microbatches == pipeline size)
        return set(range(n))
    # F^warm (F B)^(total - warm) B^warm: the forward of each steady-state pair is not a boundary
    return {i for i in range(n) if not (i >= lead - 1 and sched[i][0] == FWD and i + 1 < n and sched[i + 1][0] == BWD)}


def plan_batched(P: int, V: int, M: int, rank: int, forward_only: bool = False, sched=None):
    """Communication plan of `rank` for the batched mode: a list of (after_step, [send dst, ...], [(src, key), ...]) groups;
    after_step = -1 runs before the first step.  A group carries the tensors produced since the previous group (in
    order) and the receives that must have completed before the steps up to the next group.

    On one communicator, sends and receives between a pair of ranks match in ORDER (NCCL has no tags).  With P = 2 and
    interleaving, activations and gradients for the same neighbour share that order, and the receiver may need them in
    a different order than the sender produces them; so receives are always posted in the SENDER's order, up to the one
    needed, and tensors that arrive early wait in a dictionary.  This adds no dependency: waiting for a tensor implies
    the sender has already issued everything it sends before it."""
    sched = sched if sched is not None else build_schedule_py(P, rank, V, M, forward_only)
    last = P * V - 1
    inq, ptr = {}, {}
    for s in {(rank - 1) % P, (rank + 1) % P}:
        inq[s] = [o[1] for o in outgoing_messages(P, V, M, s, forward_only) if o is not None and o[0] == rank]
        ptr[s] = 0
    outs = outgoing_messages(P, V, M, rank, forward_only, sched)
    bnd = sorted(group_boundaries(sched, V))

    def recvs_for(steps):
        r = []
        for step in steps:
            op, c, m = sched[step]
            vs = virtual_stage(rank, c, P)
            if op == FWD and vs > 0:
                s = (rank - 1) % P
            elif op == BWD and vs < last:
                s = (rank + 1) % P
            else:
                continue
            q = inq[s]
            if (op, c, m) not in q[ptr[s]:]:
                continue                   # came in early with a previous group
            idx = q.index((op, c, m), ptr[s])
            r += [(s, q[n]) for n in range(ptr[s], idx + 1)]
            ptr[s] = idx + 1
        return r

    first = bnd[0] if bnd else len(sched) - 1
    groups = [(-1, [], recvs_for(range(0, first + 1)))]
    prev = -1
    for j, b in enumerate(bnd):
        nxt = bnd[j + 1] if j + 1 < len(bnd) else len(sched) - 1
        sends = [(i, outs[i][0]) for i in range(prev + 1, b + 1) if outs[i] is not None]
        groups.append((b, sends, recvs_for(range(b + 1, nxt + 1))))
        prev = b
    return groups


def simulate_batched(P: int, V: int, M: int, forward_only: bool = False) -> bool:
    """Deadlock check of the batched plan under the STRICTEST transport model: a group is one blocking kernel in a
    single in-order queue per rank (CUDA_DEVICE_MAX_CONNECTIONS=1), and a send completes only once the matching receive
    has started (rendezvous, no eager buffering).  True = every rank runs to completion."""
    plans = [plan_batched(P, V, M, r, forward_only) for r in range(P)]
    nsend = [dict() for _ in range(P)]
    nrecv = [dict() for _ in range(P)]
    ks = []
    for r in range(P):                     # number the messages of each directed pair
        kr = []
        for _, snd, recvs in plans[r]:
            sends = []
            for _, dst in snd:
                n = nsend[r].get(dst, 0); nsend[r][dst] = n + 1
                sends.append((r, dst, n))
            rr = []
            for s, _ in recvs:
                n = nrecv[r].get(s, 0); nrecv[r][s] = n + 1
                rr.append((s, r, n))
            kr.append((sends, rr))
        ks.append(kr)
    started_send, started_recv = set(), set()
    cur, active = [0] * P, [False] * P
    progressed = True
    while progressed:
        progressed = False
        for r in range(P):
            if cur[r] >= len(ks[r]):
                continue
            sends, recvs = ks[r][cur[r]]
            if not active[r]:
                active[r] = True
                started_send.update(sends); started_recv.update(recvs)
                progressed = True
            if all(x in started_recv for x in sends) and all(x in started_send for x in recvs):
                cur[r] += 1; active[r] = False; progressed = True
    return all(cur[r] == len(ks[r]) for r in range(P))


# ------------------------------------------------------------------------------------------
# directed p2p links
# ------------------------------------------------------------------------------------------
class P2PLinks:
    """One 2-rank group per directed neighbour link of the pipeline ring (activations: r -> next,
    gradients: r -> prev).  new_group is collective over the world, so every rank builds every link."""

    def __init__(self):
        self.act_send = self.act_recv = self.grad_send = self.grad_recv = None
        self.next_rank = self.prev_rank = None
        if not (dist.is_initialized() and mpu.model_parallel_is_initialized()):
            return
        lay = mpu._S.layout
        me = dist.get_rank()
        if lay.pp == 1:
            return
        for ranks in lay.pp_groups:
            for i, r in enumerate(ranks):
                nxt = ranks[(i + 1) % len(ranks)]
                g_act = dist.new_group([r, nxt]) if r != nxt else None
                g_grad = dist.new_group([nxt, r]) if r != nxt else None
                if me == r:
                    self.act_send, self.next_rank = g_act, nxt
                    self.grad_recv = g_grad
                if me == nxt:
                    self.act_recv, self.prev_rank = g_act, r
                    self.grad_send = g_grad


_LINKS: Optional[P2PLinks] = None


def p2p_mode() -> str:
    """`batched` (default): per schedule step ONE fused group of sends and receives (torch batch_isend_irecv = one NCCL group
    kernel) on the pipeline group's communicator, proven deadlock-free by `simulate_batched` even when every kernel of a
    rank shares one in-order hardware queue.  `links`: the unbatched dataflow mode over directed 2-rank groups with
    receives posted one step ahead; it needs a transport where a pending receive never blocks other work (gloo, or NCCL
    with enough hardware queues - with CUDA_DEVICE_MAX_CONNECTIONS=1 a receive queued ahead of the compute that feeds the
    peer stalls both ranks)."""
    m = os.environ.get("B2_PP_P2P", "batched")
    assert m in ("batched", "links"), f"B2_PP_P2P={m}: expected batched or links"
    return m


def get_links() -> P2PLinks:
    global _LINKS
    if _LINKS is None:
        _LINKS = P2PLinks()
    return _LINKS


def reset_links():
    global _LINKS
    _LINKS = None


# ------------------------------------------------------------------------------------------
# executor
# ------------------------------------------------------------------------------------------
def _custom_backward(output: torch.Tensor, grad: Optional[torch.Tensor]):
    if grad is None:
        assert output.numel() == 1
        grad = torch.ones_like(output)
    torch.autograd.backward(output, grad_tensors=grad)


def forward_backward(forward_step_func: Callable, data_iterator, model: Union[torch.nn.Module, List[torch.nn.Module]],
                     num_microbatches: int, tensor_shape: Tuple[int, ...], dtype: torch.dtype, forward_only: bool = False,
                     grad_scale: float = 1.0, collect_non_loss_data: bool = False, no_sync_func=None,
                     timers=None) -> List[Dict[str, torch.Tensor]]:
    """Runs one global batch.  `forward_step_func(data_iterator, model_chunk) -> (output, loss_func)` with
    `loss_func(output) -> (loss, {name: value})` - the Megatron callback contract (pretrain_gpt.py:103-117).
    Returns the list of per-micro-batch loss dicts on the last stage (empty elsewhere)."""
    models = model if isinstance(model, (list, tuple)) else [model]
    V = len(models)
    P = mpu.get_pipeline_model_parallel_world_size()
    rank = mpu.get_pipeline_model_parallel_rank()
    M = num_microbatches
    sched = build_schedule(P, rank, V, M, forward_only)
    batched = P > 1 and p2p_mode() == "batched"
    links = get_links() if (P > 1 and not batched) else None
    last_vs = P * V - 1
    dev = next(models[0].parameters()).device
    iters = data_iterator if isinstance(data_iterator, (list, tuple)) else [data_iterator] * V

    inputs: Dict[Tuple[int, int], Optional[torch.Tensor]] = {}
    outputs: Dict[Tuple[int, int], torch.Tensor] = {}
    pending_recv: Dict[Tuple[int, int, int], Tuple[torch.Tensor, object]] = {}
    pending_send: List[Tuple[object, torch.Tensor]] = []
    losses: List[Dict[str, torch.Tensor]] = []

    def needs_recv(op, c, m):
        vs = virtual_stage(rank, c, P)
        return (op == FWD and vs > 0) or (op == BWD and vs < last_vs)

    def post_recv(op, c, m):
        key = (op, c, m)
        if key in pending_recv or not needs_recv(op, c, m):
            return
        buf = torch.empty(tensor_shape, dtype=dtype, device=dev)
        if op == FWD:
            w = dist.irecv(buf, src=links.prev_rank, group=links.act_recv)
        else:
            w = dist.irecv(buf, src=links.next_rank, group=links.grad_recv)
        pending_recv[key] = (buf, w)

    def take_recv(op, c, m):
        buf, w = pending_recv.pop((op, c, m))
        w.wait()
        return buf

    def send(op, tensor):
        t = tensor.contiguous()
        if op == FWD:
            w = dist.isend(t, dst=links.next_rank, group=links.act_send)
        else:
            w = dist.isend(t, dst=links.prev_rank, group=links.grad_send)
        pending_send.append((w, t))

    def reap_sends(block=False):
        keep = []
        deadline = time.time() + float(os.environ.get("B2_PP_P2P_TIMEOUT_S", "300"))
        for w, t in pending_send:
            if block:
                # never hang silently: a send that does not complete means the neighbouring stage is stuck
                while t.is_cuda and not w.is_completed():
                    if time.time() > deadline:
                        raise RuntimeError("pipeline p2p send did not complete within B2_PP_P2P_TIMEOUT_S seconds "
                                           f"(pipeline rank {rank}, {len(pending_send)} sends pending): the neighbouring stage is not receiving")
                    time.sleep(0.0005)
                w.wait()
            elif not w.is_completed():
                keep.append((w, t))
        pending_send[:] = [] if block else keep

    # ---- batched mode: one fused send+recv group per boundary on the pipeline group's communicator
    arrived: Dict[Tuple[int, int, int], torch.Tensor] = {}
    produced: Dict[int, torch.Tensor] = {}
    groups_after: Dict[int, Tuple[list, list]] = {}
    send_works: List[Tuple[object, torch.Tensor]] = []
    if batched:
        pg = mpu.get_pipeline_model_parallel_group()
        pg_ranks = dist.get_process_group_ranks(pg)
        for after, snd, rcv in plan_batched(P, V, M, rank, forward_only, sched):
            groups_after[after] = (snd, rcv)

    def run_group(after):
        snd, rcv = groups_after.get(after, ((), ()))
        if not snd and not rcv:
            return
        for w, _ in send_works:            # the previous group's sends (complete by construction: see simulate_batched)
            w.wait()
        send_works.clear()
        ops, bufs, sent = [], [], []
        for step, dst in snd:
            t = produced.pop(step)
            ops.append(dist.P2POp(dist.isend, t, pg_ranks[dst], group=pg))
            sent.append(t)
        for src, key in rcv:
            buf = torch.empty(tensor_shape, dtype=dtype, device=dev)
            ops.append(dist.P2POp(dist.irecv, buf, pg_ranks[src], group=pg))
            bufs.append((key, buf))
        works = dist.batch_isend_irecv(ops)
        if len(works) == len(ops):         # one request per op (gloo): block on the receives, reap the sends later
            for w, t in zip(works[:len(sent)], sent):
                send_works.append((w, t))
            for w in works[len(sent):]:
                w.wait()
        else:                              # NCCL: one coalesced request = one fused kernel; wait() orders the compute stream after it
            for w in works:
                w.wait()
        for key, buf in bufs:
            arrived[key] = buf

    if batched:
        take_recv = lambda op, c, m: arrived.pop((op, c, m))                      # noqa: E731
        send = lambda op, tensor: produced.__setitem__(_step[0], tensor.contiguous())   # noqa: E731
        post_recv = lambda *a: None                                               # noqa: E731
        run_group(-1)
    _step = [0]
    if P > 1 and sched:
        post_recv(*sched[0])
    for i, (op, c, m) in enumerate(sched):
        _step[0] = i
        # (links mode) prefetch the receive of the NEXT step so the transfer overlaps this step's compute
        if P > 1 and i + 1 < len(sched):
            post_recv(*sched[i + 1])
        vs = virtual_stage(rank, c, P)
        mod = models[c]
        if V > 1:
            mpu.set_virtual_pipeline_model_parallel_rank(c)
        if op == FWD:
            x = None
            if vs > 0:
                x = take_recv(FWD, c, m)
                x.requires_grad_(not forward_only)
            inputs[(c, m)] = x
            unwrapped = getattr(mod, "module", mod)
            unwrapped.set_input_tensor(x)
            if timers is not None:
                timers("forward-compute", log_level=2).start()
            ctxm = torch.no_grad() if forward_only else contextlib.nullcontext()
            with ctxm:
                out, loss_func = forward_step_func(iters[c], mod)
                if vs == last_vs:
                    if collect_non_loss_data:
                        losses.append(loss_func(out, non_loss_data=True))
                        out = None
                    else:
                        loss, reduced = loss_func(out)
                        out = loss / M
                        losses.append(reduced)
            if timers is not None:
                timers("forward-compute").stop()
            if vs < last_vs:
                send(FWD, out.detach())
            if not forward_only:
                outputs[(c, m)] = out
            elif vs < last_vs:
                pass
        else:
            out = outputs.pop((c, m))
            x = inputs.pop((c, m))
            g = take_recv(BWD, c, m) if vs < last_vs else None
            is_last_bwd = all(not (o == BWD) for (o, _, _) in sched[i + 1:])
            sync_ctx = contextlib.nullcontext() if (is_last_bwd or no_sync_func is None) else no_sync_func()
            if timers is not None:
                timers("backward-compute", log_level=2).start()
            with sync_ctx:
                if vs == last_vs:
                    _custom_backward(out * grad_scale, None)
                else:
                    _custom_backward(out, g)
            if timers is not None:
                timers("backward-compute").stop()
            if vs > 0:
                send(BWD, x.grad)
        if batched:
            run_group(i)
        reap_sends()
    reap_sends(block=True)
    for w, _ in send_works:
        w.wait()
    assert not arrived and not produced, "pipeline plan left tensors undelivered"
    if V > 1:
        mpu.set_virtual_pipeline_model_parallel_rank(0)
    return losses


def get_forward_backward_func():
    """API parity with megatron.core.pipeline_parallel.get_forward_backward_func: the three reference
    schedules are the same executor driven by different generated orders."""
    return forward_backward


def get_num_microbatches(global_batch_size: int, micro_batch_size: int, data_parallel_size: int) -> int:
    """(log line: 'setting number of micro-batches to constant N', 3_...ipynb:1210)"""
    per_step = micro_batch_size * data_parallel_size
    assert global_batch_size % per_step == 0
    return global_batch_size // per_step
