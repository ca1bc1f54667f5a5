"""Table / twin tests of the host-side pieces (the reference ships no tests; these pin the behaviour its
logs show): process-group layout, pipeline schedules (C++ generator vs Python twin + deadlock simulation), Megatron
argument validation with the logged 16-GPU launch, LR schedulers, GPT dataset index builders (C++ vs Python twin),
samplers, vocab padding."""
import math

import numpy as np
import pytest
import torch

from b200dist.config import megatron_args as MA
from b200dist.data import gpt_dataset as GD
from b200dist.data.samplers import MegatronPretrainingRandomSampler, MegatronPretrainingSampler
from b200dist.optim.schedulers import OptimizerParamScheduler, WarmupDecayLR, WarmupLR
from b200dist.parallel import pipeline as PL
from b200dist.parallel.state import compute_layout


# ------------------------------------------------------------------ layout
def test_layout_16_gpus_tp2_pp4():
    lay = compute_layout(16, tp=2, pp=4)          # the worked example in Megatron's parallel_state docstring
    assert lay.dp == 2
    assert lay.dp_groups == [[0, 2], [1, 3], [4, 6], [5, 7], [8, 10], [9, 11], [12, 14], [13, 15]]
    assert lay.tp_groups == [[2 * i, 2 * i + 1] for i in range(8)]
    assert lay.pp_groups == [[0, 4, 8, 12], [1, 5, 9, 13], [2, 6, 10, 14], [3, 7, 11, 15]]
    assert lay.embedding_groups == [[g[0], g[-1]] for g in lay.pp_groups]
    assert lay.mp_groups[0] == [0, 1, 4, 5, 8, 9, 12, 13]


def test_layout_reference_run_tp4_dp4():
    lay = compute_layout(16, tp=4, pp=1)          # the logged run: TP=4, PP=1 -> DP=4 (3_...ipynb:970,1162)
    assert lay.dp == 4 and lay.tp_groups[1] == [4, 5, 6, 7] and lay.dp_groups[0] == [0, 4, 8, 12]
    assert lay.pp_groups == [[i] for i in range(16)]


@pytest.mark.parametrize("world,tp,pp", [(8, 2, 2), (8, 8, 1), (8, 1, 8), (12, 2, 3), (1, 1, 1)])
def test_layout_partitions_are_exact(world, tp, pp):
    lay = compute_layout(world, tp, pp)
    for groups, size in ((lay.tp_groups, tp), (lay.pp_groups, pp), (lay.dp_groups, lay.dp), (lay.mp_groups, tp * pp)):
        assert sorted(r for g in groups for r in g) == list(range(world)) and all(len(g) == size for g in groups)
    for r in range(world):      # any two different groups of a rank intersect only in the rank itself
        t = next(g for g in lay.tp_groups if r in g); p = next(g for g in lay.pp_groups if r in g); d = next(g for g in lay.dp_groups if r in g)
        assert set(t) & set(p) == {r} and set(t) & set(d) == {r} and set(p) & set(d) == {r}


def test_layout_rejects_indivisible():
    with pytest.raises(ValueError):
        compute_layout(6, tp=4)


# ------------------------------------------------------------------ pipeline schedule
@pytest.mark.parametrize("P,V,M", [(1, 1, 4), (2, 1, 4), (4, 1, 8), (4, 1, 2), (2, 2, 4), (4, 2, 8), (4, 3, 4), (8, 2, 16)])
def test_schedule_cpp_equals_python_and_is_deadlock_free(P, V, M):
    for r in range(P):
        a, b = PL.build_schedule(P, r, V, M), PL.build_schedule_py(P, r, V, M)
        assert a == b
        assert sum(1 for op in a if op[0] == PL.FWD) == M * V == sum(1 for op in a if op[0] == PL.BWD)
        assert PL.build_schedule(P, r, V, M, True) == PL.build_schedule_py(P, r, V, M, True)
    sim = PL.simulate(P, V, M)
    assert sim["ok"]
    if V == 1:          # 1F1B bound: stage r keeps at most P - r activations alive
        assert sim["peak_inflight"] == [min(P - r, M) for r in range(P)]


def test_interleaved_schedule_requires_divisible_microbatches():
    with pytest.raises(ValueError):
        PL.build_schedule(4, 0, 2, 6)


def test_num_microbatches():
    assert PL.get_num_microbatches(256, 4, 4) == 16
    with pytest.raises(Exception):
        PL.get_num_microbatches(10, 4, 4)


# ------------------------------------------------------------------ args
REF_LAUNCH = ("--num-layers 12 --hidden-size 768 --num-attention-heads 12 --seq-length 1024 --max-position-embeddings 1024 "
              "--micro-batch-size 12 --global-batch-size 192 --lr 0.0005 --train-iters 150000 --lr-decay-iters 150000 "
              "--lr-decay-style cosine --lr-warmup-iters 2000 --weight-decay .1 --adam-beta2 .999 --fp16 --log-interval 10 "
              "--save-interval 2000 --eval-interval 200 --eval-iters 10 --tensor-model-parallel-size 4").split()


def _validated(argv, world):
    a = MA.parse_args(argv=argv)
    a.rank, a.world_size = 0, world
    return MA.validate_args(a, {"tokenizer_type": "GPT2BPETokenizer"}, verbose=False)


def test_reference_launch_resolves_like_the_log():
    a = _validated(REF_LAUNCH, 16)
    assert a.tensor_model_parallel_size == 4 and a.pipeline_model_parallel_size == 1 and a.data_parallel_size == 4
    assert a.params_dtype == torch.float16 and a.tokenizer_type == "GPT2BPETokenizer"
    assert a.ffn_hidden_size == 3072 and a.kv_channels == 64
    assert PL.get_num_microbatches(a.global_batch_size, a.micro_batch_size, a.data_parallel_size) == 4
    assert MA.pad_vocab_size(50257, a.make_vocab_size_divisible_by, 4) == 50688          # log :1212
    cfg = MA.core_transformer_config_from_args(a)
    assert cfg.num_layers == 12 and cfg.hidden_size == 768 and cfg.num_attention_heads == 12


@pytest.mark.parametrize("extra,world", [
    (["--tensor-model-parallel-size", "3"], 16),                      # world not divisible
    (["--sequence-parallel", "--tensor-model-parallel-size", "1"], 16),  # SP needs TP > 1 -> turned off, not an error
    (["--bf16"], 16),                                                 # fp16 + bf16
    (["--num-layers-per-virtual-pipeline-stage", "2"], 16),           # interleaving without PP > 2
    (["--global-batch-size", "100"], 16),                             # not a multiple of micro * dp
])
def test_validate_args_rejects(extra, world):
    argv = [x for x in REF_LAUNCH]
    if "--tensor-model-parallel-size" in extra:
        i = argv.index("--tensor-model-parallel-size"); del argv[i:i + 2]
    if extra[0] == "--sequence-parallel":
        a = _validated(argv + extra, world)
        assert a.sequence_parallel is False
        return
    with pytest.raises((MA.ArgumentError, AssertionError, ValueError, SystemExit)):
        _validated(argv + extra, world)


def test_all_reference_flags_are_declared():
    names = set(MA.all_flag_names())
    for f in ("--num-layers --hidden-size --num-attention-heads --seq-length --max-position-embeddings --micro-batch-size "
              "--global-batch-size --lr --train-iters --lr-decay-iters --lr-decay-style --min-lr --weight-decay --lr-warmup-fraction "
              "--clip-grad --fp16 --bf16 --tensor-model-parallel-size --pipeline-model-parallel-size --sequence-parallel "
              "--data-path --vocab-file --merge-file --data-impl --split --log-interval --save-interval --eval-interval --eval-iters "
              "--save --load --distributed-backend --tokenizer-type --vocab-size --seed --use-flash-attn --recompute-activations "
              "--use-distributed-optimizer --exit-signal-handler --no-masked-softmax-fusion --no-bias-gelu-fusion").split():
        assert f in names, f


# ------------------------------------------------------------------ schedulers
class _Opt:
    def __init__(self):
        self.param_groups = [{"lr": 0.0, "weight_decay": 0.0}]


def test_megatron_cosine_schedule_points():
    s = OptimizerParamScheduler(_Opt(), 5e-4, 5e-5, 2000, 150000, "cosine", 0.1, 0.1, 150000, "constant")
    s.step(1000); assert math.isclose(s.get_lr(), 2.5e-4)
    s.step(1000); assert math.isclose(s.get_lr(), 5e-4)
    s.step(74000); assert math.isclose(s.get_lr(), 5e-5 + 0.5 * (5e-4 - 5e-5), rel_tol=1e-9)
    s.step(200000); assert s.get_lr() == 5e-5
    sd = s.state_dict()
    s2 = OptimizerParamScheduler(_Opt(), 5e-4, 5e-5, 2000, 150000, "cosine", 0.1, 0.1, 150000, "constant")
    s2.load_state_dict(sd)
    assert s2.num_steps == s.num_steps and s2.get_lr() == s.get_lr()


def test_warmup_decay_lr_like_deepspeed():
    o = _Opt()
    s = WarmupDecayLR(o, total_num_steps=100, warmup_min_lr=0.0, warmup_max_lr=1e-3, warmup_num_steps=10, warmup_type="linear")
    lrs = []
    for _ in range(100):
        s.step(); lrs.append(o.param_groups[0]["lr"])
    assert lrs[0] == pytest.approx(1e-4) or lrs[0] == pytest.approx(0.0, abs=1e-12) or lrs[0] > 0
    assert max(lrs) == pytest.approx(1e-3) and lrs[-1] < 2e-5 and all(a >= b - 1e-12 for a, b in zip(lrs[10:], lrs[11:]))
    w = WarmupLR(_Opt(), warmup_min_lr=0.0, warmup_max_lr=1e-3, warmup_num_steps=10, warmup_type="linear")
    for _ in range(50):
        w.step()
    assert w.optimizer.param_groups[0]["lr"] == pytest.approx(1e-3)


# ------------------------------------------------------------------ dataset index builders
def test_split_string():
    assert GD.get_train_valid_test_split_("969,30,1", 1000) == [0, 969, 999, 1000]
    assert GD.get_train_valid_test_split_("100", 50) == [0, 50, 50, 50]


@pytest.mark.parametrize("seq,epochs", [(16, 1), (33, 3), (7, 2)])
def test_sample_idx_cpp_equals_python(seq, epochs):
    rng = np.random.RandomState(0)
    sizes = rng.randint(1, 60, size=40).astype(np.int32)
    docs = np.arange(40, dtype=np.int32)
    doc_idx = GD.build_doc_idx(docs, epochs, np.random.RandomState(1), False)
    tpe = GD.num_tokens_per_epoch(docs, sizes)
    a = GD.build_sample_idx(sizes, doc_idx, seq, epochs, tpe)
    b = GD.build_sample_idx_py(sizes, doc_idx, seq, epochs, tpe)
    assert a.shape == b.shape == ((epochs * tpe - 1) // seq + 1, 2) and np.array_equal(a, b)
    # every sample spans exactly seq+1 tokens
    flat = np.concatenate([np.full(sizes[d], d) for d in doc_idx])
    assert len(flat) >= (len(a) - 1) * seq + 1


def test_blending_indices_follow_weights():
    idx, sample = GD.build_blending_indices(np.array([0.7, 0.2, 0.1]), 1000)
    counts = np.bincount(idx, minlength=3)
    assert abs(counts[0] - 700) <= 1 and abs(counts[1] - 200) <= 1 and abs(counts[2] - 100) <= 1
    for d in range(3):
        assert np.array_equal(sample[idx == d], np.arange(counts[d]))


def test_num_epochs():
    assert GD.num_epochs_for(1000, 10, 99) == 1 and GD.num_epochs_for(1000, 10, 100) == 2


def test_pretraining_sampler_shards_and_resumes():
    full = [list(b) for b in MegatronPretrainingSampler(64, 0, 4, 0, 1)]
    r0 = [list(b) for b in MegatronPretrainingSampler(64, 0, 2, 0, 2)]
    r1 = [list(b) for b in MegatronPretrainingSampler(64, 0, 2, 1, 2)]
    assert [a + b for a, b in zip(r0, r1)] == full
    resumed = [list(b) for b in MegatronPretrainingSampler(64, 16, 2, 0, 2)]
    assert resumed == r0[4:]
    rnd0 = [list(b) for b in MegatronPretrainingRandomSampler(64, 0, 2, 0, 2)]
    rnd1 = [list(b) for b in MegatronPretrainingRandomSampler(64, 0, 2, 1, 2)]
    assert sorted(i for b in rnd0 for i in b) == list(range(32)) and sorted(i for b in rnd1 for i in b) == list(range(32, 64))
    assert [list(b) for b in MegatronPretrainingRandomSampler(64, 16, 2, 0, 2)] == rnd0[4:]


# ---------------------------------------------------------------------------------------------------------------
# batched pipeline p2p plan
# ---------------------------------------------------------------------------------------------------------------
def test_batched_p2p_plan_is_deadlock_free_on_an_in_order_rendezvous_transport():
    """every (P, V, M): one blocking kernel per group, one queue per rank, sends complete only against a started receive"""
    from b200dist.parallel import pipeline as pp
    n = 0
    for P in (2, 3, 4, 8):
        for V in (1, 2, 3, 4):
            for M in range(1, 25):
                if V > 1 and M % P:
                    continue
                for fo in (False, True):
                    assert pp.simulate_batched(P, V, M, fo), (P, V, M, fo)
                    n += 1
    assert n > 300


def test_batched_p2p_plan_delivers_every_tensor_once_and_in_sender_order():
    from b200dist.parallel import pipeline as pp
    for P, V, M in ((2, 1, 4), (2, 2, 4), (4, 2, 8), (3, 3, 6), (2, 3, 2)):
        plans = [pp.plan_batched(P, V, M, r) for r in range(P)]
        for r in range(P):
            sched = pp.build_schedule_py(P, r, V, M)
            need = [(op, c, m) for (op, c, m) in sched
                    if (op == pp.FWD and pp.virtual_stage(r, c, P) > 0) or (op == pp.BWD and pp.virtual_stage(r, c, P) < P * V - 1)]
            got = [key for _, _, rcv in plans[r] for _, key in rcv]
            assert sorted(got) == sorted(need)
            # a tensor is received in a group that runs before the step that consumes it
            pos = {st: i for i, st in enumerate(sched)}
            for after, _, rcv in plans[r]:
                for _, key in rcv:
                    assert pos[key] > after
            # per source, receive order == that source's send order
            for src in {(r - 1) % P, (r + 1) % P}:
                sent = [o[1] for o in pp.outgoing_messages(P, V, M, src) if o is not None and o[0] == r]
                recv = [key for _, _, rcv in plans[r] for s, key in rcv if s == src]
                assert recv == sent


# ---------------------------------------------------------------------------------------------------------------
# sentence / block sample mappings (helpers.cpp build_mapping, build_blocks_mapping)
# ---------------------------------------------------------------------------------------------------------------
def _sentence_corpus(seed=0, n_docs=40):
    rs = np.random.RandomState(seed)
    counts = rs.randint(1, 9, size=n_docs)
    docs = np.concatenate([[0], np.cumsum(counts)]).astype(np.int64)
    sizes = rs.randint(3, 60, size=int(docs[-1])).astype(np.int32)
    sizes[int(docs[5])] = 600                     # a document with an over-long sentence is skipped
    return docs, sizes


def test_build_mapping_matches_python_twin_and_is_a_seeded_permutation():
    from b200dist.data import dataset_utils as du
    docs, sizes = _sentence_corpus()
    for prob in (0.0, 0.1, 0.5):
        raw = du.build_mapping(docs, sizes, 3, 10 ** 6, 128, prob, 1234, shuffle=False)
        twin = du.build_mapping_py(docs, sizes, 3, 10 ** 6, 128, prob, 1234)
        assert raw.shape == twin.shape and (raw == twin).all()
        sh = du.build_mapping(docs, sizes, 3, 10 ** 6, 128, prob, 1234)
        assert sorted(map(tuple, sh)) == sorted(map(tuple, raw)) and not (sh == raw).all()
        assert (sh == du.build_mapping(docs, sizes, 3, 10 ** 6, 128, prob, 1234)).all()       # deterministic per seed
    for a, b, t in raw:
        d = np.searchsorted(docs, a, side="right") - 1
        assert docs[d] <= a < b <= docs[d + 1] and d != 5 and 2 <= t <= 128
    # the epoch loop stops once max_num_samples is reached
    one = du.build_mapping(docs, sizes, 1, 10 ** 6, 128, 0.0, 1, shuffle=False)
    capped = du.build_mapping(docs, sizes, 5, len(one), 128, 0.0, 1, shuffle=False)
    assert len(capped) == len(one)


def test_build_blocks_mapping_matches_python_twin():
    from b200dist.data import dataset_utils as du
    docs, sizes = _sentence_corpus(1)
    titles = np.random.RandomState(2).randint(2, 12, size=len(docs) - 1).astype(np.int32)
    for one_sent in (False, True):
        raw = du.build_blocks_mapping(docs, sizes, titles, 2, 10 ** 6, 96, 7, one_sent, shuffle=False)
        twin = du.build_blocks_mapping_py(docs, sizes, titles, 2, 10 ** 6, 96, 7, one_sent)
        assert raw.shape == twin.shape and (raw == twin).all()
        sh = du.build_blocks_mapping(docs, sizes, titles, 2, 10 ** 6, 96, 7, one_sent)
        assert sorted(map(tuple, sh)) == sorted(map(tuple, raw))
    half = len(raw) // 2
    assert (raw[:half, 3] == np.arange(half)).all() and raw[half, 3] == 0                   # block ids restart every epoch
    for a, b, d, _ in raw:
        assert docs[d] <= a < b <= docs[d + 1]


def test_recompute_plan_methods():
    from b200dist.models.gpt import recompute_plan
    assert recompute_plan(5, None, 3) == [(0, 1, True), (1, 2, True), (2, 3, True), (3, 4, True), (4, 5, True)]
    assert recompute_plan(5, "uniform", 2) == [(0, 2, True), (2, 4, True), (4, 5, True)]
    assert recompute_plan(4, "block", 1) == [(0, 1, True), (1, 2, False), (2, 3, False), (3, 4, False)]
    with pytest.raises(ValueError):
        recompute_plan(4, "other", 1)


def test_microbatch_calculators_constant_and_rampup():
    from types import SimpleNamespace
    from b200dist.utils import microbatches as mb
    a = SimpleNamespace(rampup_batch_size=None, global_batch_size=192, micro_batch_size=12, data_parallel_size=4, rank=1)
    mb.build_num_microbatches_calculator(a)
    assert mb.get_num_microbatches() == 4 and mb.get_current_global_batch_size() == 192
    a.rampup_batch_size = ["32", "32", "1000"]          # 32 -> 192 in 5 increments of 32 over 1000 samples (200 samples each)
    a.micro_batch_size, a.data_parallel_size = 4, 2
    mb.build_num_microbatches_calculator(a)
    seen = []
    for consumed in (0, 199, 200, 450, 999, 1000, 1001, 5000):
        mb.update_num_microbatches(consumed)
        seen.append((mb.get_current_global_batch_size(), mb.get_num_microbatches()))
    assert seen == [(32, 4), (32, 4), (64, 8), (96, 12), (160, 20), (192, 24), (192, 24), (192, 24)]
    a.micro_batch_size = 5                               # 32 is not a multiple of 5 * 2: caught by the consistency check
    mb.build_num_microbatches_calculator(a)
    with pytest.raises(AssertionError):
        mb.update_num_microbatches(0, consistency_check=True)


def test_hf_schedulers_match_transformers():
    transformers = pytest.importorskip("transformers")
    from b200dist.optim.schedulers import HFScheduler
    for name in ("linear", "cosine", "constant", "constant_with_warmup", "inverse_sqrt"):
        w = torch.nn.Parameter(torch.zeros(1))
        ref_opt = torch.optim.SGD([w], lr=0.5)
        ref = transformers.get_scheduler(name, ref_opt, num_warmup_steps=3, num_training_steps=20)
        class Opt:
            param_groups = [{"lr": 0.5}]
        ours = HFScheduler(Opt, name, 3, 20)
        for _ in range(20):
            assert abs(Opt.param_groups[0]["lr"] - ref_opt.param_groups[0]["lr"]) < 1e-9, (name, ours.last_step)
            ref_opt.step(); ref.step(); ours.step()
        sd = ours.state_dict()
        again = HFScheduler(Opt, name, 3, 20)
        again.load_state_dict(sd)
        assert again.get_last_lr() == ours.get_last_lr()


def test_prepare_oxford_pet_split(tmp_path):
    import importlib.util
    import os
    from PIL import Image
    spec = importlib.util.spec_from_file_location("prep", os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools", "prepare_oxford_pet.py"))
    prep = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(prep)
    src = tmp_path / "images"
    src.mkdir()
    for cls in ("Abyssinian", "american_bulldog"):
        for i in range(1, 21):
            Image.new("RGB", (8, 8), (i * 10, 0, 0)).save(src / f"{cls}_{i}.jpg")
    (src / "american_bulldog_99.jpg").write_bytes(b"not a jpeg")                 # undecodable -> dropped
    r = prep.prepare(str(src), str(tmp_path / "out"))
    assert r["classes"] == 2 and r["train"] == 15 + 16 and r["val"] == 1 + 2 and r["test"] == 3 + 2 and "american_bulldog_99.jpg" in r["dropped"]
    assert "Abyssinian_5.jpg" in r["dropped"]                                    # the notebook's list of corrupt files
    assert sorted(os.listdir(tmp_path / "out")) == ["test", "train", "val"]
    assert sorted(os.listdir(tmp_path / "out" / "train")) == ["Abyssinian", "american_bulldog"]
    again = prep.prepare(str(src), str(tmp_path / "out2"))
    assert sorted(os.listdir(tmp_path / "out" / "val" / "Abyssinian")) == sorted(os.listdir(tmp_path / "out2" / "val" / "Abyssinian"))


def test_tensor_parallel_checkpoint_repeats_dropout_masks():
    from b200dist.parallel import tensor as tpo
    torch.manual_seed(0)
    tpo.get_cuda_rng_tracker().reset()
    tpo.get_cuda_rng_tracker().add("model-parallel-rng", 4321)
    lin = torch.nn.Linear(16, 16)

    def block(x, scale):
        h = torch.nn.functional.dropout(lin(x), 0.5, training=True)            # default stream
        with tpo.get_cuda_rng_tracker().fork():
            h = torch.nn.functional.dropout(h, 0.3, training=True)             # tracker stream
        return h * scale

    x = torch.randn(4, 16, requires_grad=True)
    st_cpu, st_tr = torch.get_rng_state(), tpo.get_cuda_rng_tracker().get_states()
    y = block(x, 2.0)
    y.sum().backward()
    want = (y.detach().clone(), x.grad.clone(), lin.weight.grad.clone())
    x.grad = None; lin.zero_grad()
    torch.set_rng_state(st_cpu); tpo.get_cuda_rng_tracker().set_states(st_tr)
    y2 = tpo.checkpoint(block, False, x, 2.0)
    torch.nn.functional.dropout(torch.ones(8), 0.5, training=True)           # RNG use between forward and backward must not matter
    y2.sum().backward()
    assert torch.equal(y2, want[0]) and torch.equal(x.grad, want[1]) and torch.equal(lin.weight.grad, want[2])


def _state_extras_worker(rank, world):
    import torch.distributed as dist
    from b200dist.parallel import state as mpu
    mpu.initialize_model_parallel(1, 2, None, 1)                  # PP=2 with the split after stage 0; DP groups are singletons at world 2
    g = mpu.get_data_parallel_group_gloo()
    before, after, at = mpu.is_pipeline_stage_before_split(), mpu.is_pipeline_stage_after_split(), mpu.is_pipeline_stage_at_split()
    buf = mpu.get_global_memory_buffer()
    a = buf.get_tensor((4, 8), torch.float32, "mpu", device="cpu")
    b = buf.get_tensor((2, 8), torch.float32, "mpu", device="cpu")
    c = buf.get_tensor((8, 8), torch.float32, "mpu", device="cpu")
    return (g is not None, before, after, at, a.data_ptr() == b.data_ptr(), tuple(c.shape), mpu.get_pipeline_model_parallel_split_rank())


def test_parallel_state_split_rank_gloo_group_and_memory_buffer():
    from tests._dist_util import run_distributed
    outs = run_distributed(_state_extras_worker, 2)
    assert outs[0][1:4] == (True, False, True) and outs[1][1:4] == (False, True, False)
    assert all(o[0] and o[4] and o[5] == (8, 8) and o[6] == 1 for o in outs)


def test_indexed_dataset_formats_mmap_lazy_cached_infer(tmp_path):
    from b200dist.data import indexed_dataset as idx
    rs = np.random.RandomState(0)
    docs = [[rs.randint(0, 1000, size=rs.randint(1, 20)) for _ in range(rs.randint(1, 4))] for _ in range(12)]
    for impl in ("mmap", "cached"):
        prefix = str(tmp_path / impl)
        b = idx.make_builder(idx.data_file_path(prefix), impl, dtype=np.uint16)
        for d in docs:
            for sent in d:
                b.add_item(sent)
            b.end_document()
        b.finalize(idx.index_file_path(prefix))
        assert idx.infer_dataset_impl(prefix) == impl
    flat = [s for d in docs for s in d]
    views = {"mmap": idx.make_dataset(str(tmp_path / "mmap"), "mmap"), "infer-mmap": idx.make_dataset(str(tmp_path / "mmap"), "infer"),
             "lazy": idx.make_dataset(str(tmp_path / "cached"), "lazy"), "cached": idx.make_dataset(str(tmp_path / "cached"), "infer")}
    assert isinstance(views["cached"], idx.IndexedCachedDataset) and isinstance(views["infer-mmap"], idx.MMapIndexedDataset)
    views["cached"].prefetch(range(0, len(flat), 2))
    for name, ds in views.items():
        assert len(ds) == len(flat) and list(ds.sizes) == [len(s) for s in flat], name
        assert list(ds.doc_idx) == list(np.cumsum([0] + [len(d) for d in docs])), name
        for i, s in enumerate(flat):
            assert (ds[i] == s).all() and (ds.get(i, 0, 1) == s[:1]).all(), (name, i)
            if len(s) > 2:
                assert (ds.get(i, 1, len(s) - 2) == s[1:-1]).all()
    # the GPT dataset builds the same samples from either format
    from b200dist.data.gpt_dataset import build_train_valid_test_datasets
    out = []
    for impl, pref in (("mmap", "mmap"), ("lazy", "cached")):
        tr, va, te = build_train_valid_test_datasets([str(tmp_path / pref)], impl, "8,1,1", [20, 4, 4], 16, 1234, True)
        out.append([tr[i]["text"].tolist() for i in range(5)])
    assert out[0] == out[1]


def test_dynamic_loss_scale_reproduces_the_reference_trace():
    """3_training_megatron-lm.ipynb:4347-4348: initial scale 2^32, hysteresis 2; iterations 1-10 all overflow -> the log shows
    8388608 (2^23) at iteration 10; 8 more skipped iterations by iteration 20 -> 32768 (2^15); then growth after `window` clean steps"""
    from b200dist.optim.mixed import DynamicGradScaler
    s = DynamicGradScaler(initial_scale=2.0 ** 32, min_scale=1.0, growth_factor=2.0, backoff_factor=0.5, growth_interval=1000, hysteresis=2)
    for _ in range(10):
        s.update(True)
    assert s.scale == 8388608.0
    for _ in range(8):
        s.update(True)
    assert s.scale == 32768.0
    s.update(False); s.update(False)
    assert s.scale == 32768.0
    sd = s.state_dict()
    t = DynamicGradScaler(initial_scale=1.0, growth_interval=1000, hysteresis=2)
    t.load_state_dict(sd)
    for _ in range(998):
        t.update(False)
    assert t.scale == 65536.0                       # 1000 clean steps in total: one growth step
    t.update(True)                                  # hysteresis was re-armed by the growth: the first overflow does not back off
    assert t.scale == 65536.0
    t.update(True)
    assert t.scale == 32768.0
    m = DynamicGradScaler(initial_scale=4.0, min_scale=2.0, hysteresis=1)
    for _ in range(5):
        m.update(True)
    assert m.scale == 2.0                           # floor at --min-loss-scale


def test_merge_datasets_tool(tmp_path):
    import importlib.util
    import os
    from b200dist.data import indexed_dataset as idx
    spec = importlib.util.spec_from_file_location("merge_datasets", os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools", "merge_datasets.py"))
    tool = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(tool)
    rs = np.random.RandomState(1)
    all_docs = []
    for k, impl in enumerate(("mmap", "cached")):
        prefix = str(tmp_path / f"part{k}")
        b = idx.make_builder(idx.data_file_path(prefix), impl, dtype=np.uint16)
        for _ in range(5):
            doc = [rs.randint(0, 500, size=rs.randint(1, 9)) for _ in range(rs.randint(1, 3))]
            for s in doc:
                b.add_item(s)
            b.end_document()
            all_docs.append(doc)
        b.finalize(idx.index_file_path(prefix))
    n_items, n_docs = tool.merge([str(tmp_path / "part0"), str(tmp_path / "part1")], str(tmp_path / "merged"))
    ds = idx.make_dataset(str(tmp_path / "merged"), "infer")
    flat = [s for d in all_docs for s in d]
    assert n_docs == 10 and n_items == len(flat) == len(ds) and ds[0].dtype == np.uint16
    assert all((ds[i] == s).all() for i, s in enumerate(flat))
    assert list(ds.doc_idx) == list(np.cumsum([0] + [len(d) for d in all_docs]))


def test_gpt_dataset_index_cache_goes_to_data_cache_path_on_a_read_only_data_dir(tmp_path):
    """--data-cache-path is where a missing index cache is WRITTEN (reference gpt_dataset.py:345-366): a read-only data
    channel keeps working; a second build finds the files there and writes nothing"""
    import os, stat
    from b200dist.data import indexed_dataset as idx
    from b200dist.data.gpt_dataset import build_train_valid_test_datasets
    data = tmp_path / "data"; cache = tmp_path / "cache"
    data.mkdir()
    rs = np.random.RandomState(1)
    b = idx.make_builder(idx.data_file_path(str(data / "corpus")), "mmap", dtype=np.uint16)
    for _ in range(30):
        b.add_item(rs.randint(0, 500, size=rs.randint(5, 40))); b.end_document()
    b.finalize(idx.index_file_path(str(data / "corpus")))
    os.chmod(data, stat.S_IRUSR | stat.S_IXUSR)          # read-only channel
    try:
        tr, va, te = build_train_valid_test_datasets([str(data / "corpus")], "mmap", "8,1,1", [12, 2, 2], 16, 1234, True, data_cache_path=str(cache))
        assert not (data / "index-cache").exists()
        files = sorted(os.listdir(cache))
        assert any(f.endswith("_doc_idx.npy") for f in files) and any(f.endswith(".dsc") for f in files)
        first = [tr[i]["text"].tolist() for i in range(4)]
        mtimes = {f: os.path.getmtime(cache / f) for f in files}
        tr2, _, _ = build_train_valid_test_datasets([str(data / "corpus")], "mmap", "8,1,1", [12, 2, 2], 16, 1234, True, data_cache_path=str(cache))
        assert sorted(os.listdir(cache)) == files and all(os.path.getmtime(cache / f) == t for f, t in mtimes.items())
        assert [tr2[i]["text"].tolist() for i in range(4)] == first
    finally:
        os.chmod(data, stat.S_IRWXU)


def test_gemm_plan_split_k_cost_model():
    """csrc/gemm.cu's host-side planning (tile width, CTA pairs, split-K for fp32-accumulate outputs) through b2_gemm_plan: the split it
    picks is the smallest within 4 % of the best of its own cost model, never leaves a split without k-blocks, and the weight-gradient
    shapes of the GPT-2 small benchmark get the splits that were measured on the B200 (profiles/r2_kernel_check_gemm_v3_wgrad.log)"""
    import random
    from b200dist.ops import native as N
    if not N.available():
        pytest.skip("libb200dist.so not built")

    def model(tiles, num_kb, slots):
        def cost(s):
            kbs = -(-num_kb // s)
            items = tiles * -(-num_kb // kbs)
            return -(-items // slots) * max(kbs + 4, 32)
        cands = [1] + [s for s in range(2, 33) if num_kb // s >= 4]
        best = min(cost(s) for s in cands)
        return next(s for s in cands if cost(s) * 100 <= best * 104)
    rnd = random.Random(0)
    shapes = [(2304, 768, 24576), (768, 768, 24576), (3072, 768, 24576), (768, 3072, 24576), (32768, 768, 24576), (1152, 768, 24576)]
    shapes += [(rnd.randrange(1, 64) * 64, rnd.randrange(1, 64) * 64, rnd.randrange(1, 600) * 64) for _ in range(200)]
    for (M, Nn, K) in shapes:
        for sms in (148, 124):
            g = N.gemm_plan(M, Nn, K, out_fp32=True, accumulate=True, sms=sms)
            tiles, slots = g["tiles_m"] * g["tiles_n"], max(1, sms // 2 if g["pair"] else sms)
            want = model(tiles, g["num_kb"], slots)
            kbs = -(-g["num_kb"] // want)
            assert g["kb_per_split"] == kbs and g["split_k"] == -(-g["num_kb"] // kbs), (M, Nn, K, sms, g, want)
            assert (g["split_k"] - 1) * g["kb_per_split"] < g["num_kb"] <= g["split_k"] * g["kb_per_split"]
            assert g["tiles_m"] == -(-M // (256 if g["pair"] else 128)) and g["tiles_n"] == -(-Nn // g["bn"])
    assert N.gemm_plan(2304, 768, 24576, True, True)["split_k"] == 5           # 27 pair tiles on 74 pairs: two rounds of 77 k-blocks
    assert N.gemm_plan(3072, 768, 24576, True, True)["split_k"] == 2           # 36 pair tiles: one round
    assert N.gemm_plan(768, 768, 24576, True, True)["split_k"] == 8
    for (M, Nn, K) in [(12288, 2304, 768), (12288, 768, 3072), (4096, 11008, 4096)]:      # forward / dgrad shapes never split
        g = N.gemm_plan(M, Nn, K)
        assert g["split_k"] == 1 and g["pair"] == 1 and g["bn"] == 256
    assert N.gemm_plan(512, 128, 1024)["bn"] == 128


def test_reset_position_ids_matches_the_reference_loop():
    """--reset-position-ids: vectorised restart of the positions after each EOD against get_ltor_masks_and_position_ids' per-document loop"""
    import importlib.util
    import os
    spec = importlib.util.spec_from_file_location("pretrain_gpt_entry", os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "entrypoints", "pretrain_gpt.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    g = torch.Generator().manual_seed(0)
    for _ in range(20):
        t = torch.randint(0, 6, (4, 33), generator=g)
        got = mod.reset_position_ids_after_eod(t, 5)
        ref = torch.arange(33).unsqueeze(0).repeat(4, 1)
        for b in range(4):
            prev = 0
            for i in (t[b] == 5).nonzero().flatten().tolist():
                ref[b, i + 1:] -= (i + 1 - prev)
                prev = i + 1
        assert torch.equal(got, ref)
