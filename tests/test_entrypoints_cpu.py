"""Entry-point level tests on CPU (the reference's only "test strategy" is running its notebooks' launch lines and
reading the logs): Megatron GPT pretrain with bit-exact checkpoint resume and the log-line format,
MNIST DDP script, Oxford-Pet Swin script on a synthetic image folder."""
import os
import re
import subprocess
import sys

import pytest


def _free_port():
    import socket
    with socket.socket() as so:
        so.bind(("127.0.0.1", 0))
        return so.getsockname()[1]


ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GPT = ("--num-layers 2 --hidden-size 64 --num-attention-heads 4 --seq-length 32 --max-position-embeddings 32 --micro-batch-size 2 "
       "--global-batch-size 4 --lr 1e-3 --lr-decay-style cosine --lr-warmup-iters 2 --weight-decay 0.1 --log-interval 1 "
       "--eval-interval 100 --eval-iters 1 --tokenizer-type NullTokenizer --vocab-size 255 --mock-data --distributed-backend gloo "
       "--seed 7 --train-iters 6 --lr-decay-iters 6").split()


def _run(script, argv, nproc=0, timeout=600):
    if nproc:
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}", "--master-addr", "127.0.0.1",
               "--master-port", str(_free_port()), os.path.join(ROOT, "entrypoints", script), *argv]
    else:
        cmd = [sys.executable, os.path.join(ROOT, "entrypoints", script), *argv]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, env={**os.environ, "OMP_NUM_THREADS": "2", "MKL_CBWR": "AVX2"})
    assert r.returncode == 0, (r.stdout[-3000:] + r.stderr[-3000:])
    return r.stdout + r.stderr


class _Losses(dict):
    """{iteration: printed 'lm loss'}.  Two runs are EQUAL when they logged the same iterations and every loss agrees to the printed
    precision (7 significant digits, +-2 in the last one): the runs compared here are separate processes, and the CPU BLAS under them is
    reproducible only up to how the allocator happened to align its buffers."""

    def __eq__(self, other):
        if not isinstance(other, dict) or sorted(self) != sorted(other):
            return False
        return all(abs(float(self[k]) - float(other[k])) <= 2.5e-6 * max(abs(float(self[k])), 1e-30) for k in self)

    def __ne__(self, other):
        return not self.__eq__(other)

    __hash__ = None


def _close(a, b):
    """two printed losses agree to the printed precision (see _Losses)"""
    return abs(float(a) - float(b)) <= 2.5e-6 * max(abs(float(a)), 1e-30)


def _losses(out):
    return _Losses({int(m.group(1)): m.group(2) for m in re.finditer(r"iteration\s+(\d+)/\s*\d+ \|.*?lm loss: ([0-9.E+-]+)", out)})


def test_gpt_pretrain_log_format_and_bit_exact_resume(tmp_path):
    ck = str(tmp_path / "ck")
    out = _run("pretrain_gpt.py", GPT + ["--save", ck, "--save-interval", "3"])
    full = _losses(out)
    assert sorted(full) == [1, 2, 3, 4, 5, 6]
    # the Megatron training-log line the notebook shows (3_...ipynb:1393)
    line = next(l for l in out.splitlines() if "iteration        1/" in l)
    for field in ("consumed samples:", "elapsed time per iteration (ms):", "learning rate:", "global batch size:", "lm loss:",
                  "loss scale:", "grad norm:", "number of skipped iterations:", "number of nan iterations:"):
        assert field in line
    assert os.path.isfile(os.path.join(ck, "iter_0000003", "mp_rank_00", "model_optim_rng.pt"))
    assert open(os.path.join(ck, "latest_checkpointed_iteration.txt")).read().strip() == "6"
    with open(os.path.join(ck, "latest_checkpointed_iteration.txt"), "w") as f:
        f.write("3")
    resumed = _losses(_run("pretrain_gpt.py", GPT + ["--load", ck]))
    assert sorted(resumed) == [4, 5, 6] and all(_close(resumed[i], full[i]) for i in (4, 5, 6))      # dropout masks included


def test_gpt_pretrain_tp2_checkpoint_names(tmp_path):
    ck = str(tmp_path / "ck")
    argv = [a for a in GPT]
    argv[argv.index("--train-iters") + 1] = "2"
    _run("pretrain_gpt.py", argv + ["--tensor-model-parallel-size", "2", "--sequence-parallel", "--save", ck, "--save-interval", "2"], nproc=2)
    assert sorted(os.listdir(os.path.join(ck, "iter_0000002"))) == ["mp_rank_00", "mp_rank_01"]


def test_mnist_ddp_script(tmp_path):
    out = _run("pytorch_mnist_ddp.py", ["--epochs", "1", "--batch-size", "32", "--test-batch-size", "64", "--synthetic", "--synthetic-size", "256",
                                        "--model-dir", str(tmp_path), "--backend", "gloo", "--log-interval", "2", "--save-model"], nproc=2)
    assert "Train Epoch: 1" in out and "Test set: Average loss:" in out
    assert os.path.isfile(tmp_path / "mnist_cnn.pt")


def test_oxford_swin_script_synthetic(tmp_path):
    out = _run("pytorch_oxford_ddp.py", ["--model_name", "swin_t", "--num-classes", "37", "--num-epochs", "1", "--batch-size", "2",
                                         "--test-batch-size", "2", "--synthetic", "--synthetic-train-size", "8", "--synthetic-val-size", "4",
                                         "--height", "64", "--width", "64", "--backend", "gloo", "--no-pretrained", "--num-workers", "0",
                                         "--model-dir", str(tmp_path), "--print-freq", "1", "--max-steps", "2"], nproc=2, timeout=900)
    assert "Epoch: [0]" in out or "Epoch: [1]" in out
    assert "Test_Prec@1=" in out and "Test_Loss=" in out


def test_gpt_standalone_embedding_stage_matches_single_process():
    """--standalone-embedding-stage (arguments.py:1054): pipeline stage 0 holds only the embedding, 3 layers on the other 3 stages of PP=4"""
    base = [a for a in GPT] + ["--hidden-dropout", "0", "--attention-dropout", "0"]
    base[base.index("--num-layers") + 1] = "3"
    base[base.index("--global-batch-size") + 1] = "8"
    base[base.index("--train-iters") + 1] = "3"
    base[base.index("--lr-decay-iters") + 1] = "3"
    single = _losses(_run("pretrain_gpt.py", base))
    staged = _losses(_run("pretrain_gpt.py", base + ["--pipeline-model-parallel-size", "4", "--standalone-embedding-stage"], nproc=4))
    assert sorted(single) == [1, 2, 3] and single == staged, (single, staged)


def test_mnist_script_reads_raw_idx_files(tmp_path):
    """--data-path with torchvision's on-disk layout (<root>/MNIST/raw/*-idx?-ubyte, plain and gzipped): the script trains on the files,
    not on the synthetic stand-in (reference: datasets.MNIST(args.data_path, ...), pytorch_mnist_ddp.py:216-228)"""
    import gzip
    import numpy as np
    import torch
    from b200dist.data.mnist import MNISTTensors, write_idx
    rng = np.random.default_rng(0)
    raw = tmp_path / "MNIST" / "raw"
    tr_x, tr_y = rng.integers(0, 255, (96, 28, 28), dtype=np.uint8), rng.integers(0, 10, (96,), dtype=np.uint8)
    te_x, te_y = rng.integers(0, 255, (32, 28, 28), dtype=np.uint8), rng.integers(0, 10, (32,), dtype=np.uint8)
    write_idx(str(raw / "train-images-idx3-ubyte"), tr_x)
    write_idx(str(raw / "train-labels-idx1-ubyte"), tr_y)
    for name, arr in (("t10k-images-idx3-ubyte", te_x), ("t10k-labels-idx1-ubyte", te_y)):      # the test split gzipped
        write_idx(str(raw / name), arr)
        with open(raw / name, "rb") as f, gzip.open(str(raw / name) + ".gz", "wb") as gz:
            gz.write(f.read())
        os.remove(raw / name)
    ds = MNISTTensors(str(tmp_path), train=True)
    assert not ds.synthetic and len(ds) == 96 and torch.equal(ds.labels, torch.from_numpy(tr_y.astype(np.int64)))
    assert torch.allclose(ds.images[:, 0], (torch.from_numpy(tr_x).float() / 255 - 0.1307) / 0.3081)
    te = MNISTTensors(str(tmp_path), train=False)
    assert not te.synthetic and len(te) == 32 and torch.equal(te.labels, torch.from_numpy(te_y.astype(np.int64)))
    out = _run("pytorch_mnist.py", ["--data-path", str(tmp_path), "--epochs", "1", "--batch-size", "32", "--test-batch-size", "16", "--log-interval", "1"])
    assert "Train Epoch: 1 [64/96" in out and "Test set: Average loss:" in out and "/32 (" in out


def test_oxford_script_real_image_folder_host_and_device_augmentation(tmp_path):
    """<data-dir>/{train,val}/<class>/*.jpg through the script: once with the reference's host-side augmentation pipeline and once
    with --gpu-augment (workers decode + resize, uint8 batches, random ops on the training device)"""
    cv2 = pytest.importorskip("cv2")
    import numpy as np
    rng = np.random.default_rng(0)
    for split, n in (("train", 6), ("val", 2)):
        for cls in ("abyssinian", "beagle"):
            d = tmp_path / "data" / split / cls
            d.mkdir(parents=True)
            for i in range(n):
                cv2.imwrite(str(d / f"{cls}_{i}.jpg"), rng.integers(0, 255, (rng.integers(60, 90), rng.integers(60, 90), 3), dtype=np.uint8))
    common = ["--model_name", "swin_t", "--num-classes", "2", "--num-epochs", "1", "--batch-size", "4", "--test-batch-size", "2", "--height", "64",
              "--width", "64", "--backend", "gloo", "--no-pretrained", "--num-workers", "0", "--data-dir", str(tmp_path / "data"), "--print-freq", "1",
              "--log-interval", "1", "--max-steps", "2"]
    for extra in ([], ["--gpu-augment"]):
        out = _run("pytorch_oxford_ddp.py", common + ["--model-dir", str(tmp_path / ("m" + str(len(extra))))] + extra, timeout=900)
        assert "Train_Loss=" in out and "Test_Prec@1=" in out, out[-2000:]
        assert os.path.isfile(tmp_path / ("m" + str(len(extra))) / "checkpoint.pth")


def test_swin_b_matches_torchvision_structure():
    """same parameter names, shapes and count as torchvision's swin_b (the reference builds it by name, train_util:18-41)."""
    tv = pytest.importorskip("torchvision")
    import torch
    from b200dist.models import swin
    with torch.device("meta"):
        ref = tv.models.swin_b(weights=None)
        ours = swin.swin_b()
    a = {k: tuple(v.shape) for k, v in ref.state_dict().items()}
    b = {k: tuple(v.shape) for k, v in ours.state_dict().items()}
    assert a == b and sum(p.numel() for p in ours.parameters()) == 87_768_224


def test_gpt_pretrain_tp2_pp2_interleaved_matches_single_process():
    """4 gloo ranks, TP=2 (+SP) x PP=2 with 2 model chunks per rank: same loss curve as one process (each model chunk draws
    the same micro-batch sequence from its own data iterator)."""
    base = [a for a in GPT]
    base[base.index("--num-layers") + 1] = "8"
    base[base.index("--global-batch-size") + 1] = "8"
    base[base.index("--train-iters") + 1] = "3"
    base += ["--hidden-dropout", "0", "--attention-dropout", "0", "--eval-iters", "0"]
    single = _losses(_run("pretrain_gpt.py", base))
    par = _losses(_run("pretrain_gpt.py", base + ["--tensor-model-parallel-size", "2", "--sequence-parallel", "--pipeline-model-parallel-size", "2",
                                                  "--num-layers-per-virtual-pipeline-stage", "2"], nproc=4, timeout=900))
    assert sorted(single) == [1, 2, 3] == sorted(par)
    assert all(abs(float(single[i]) - float(par[i])) < 2e-4 for i in (1, 2, 3)), (single, par)


def test_gpt_full_recompute_is_bit_identical_with_dropout():
    """--recompute-granularity full re-runs each layer in backward with the dropout seeds of its first run."""
    base = [a for a in GPT]
    base[base.index("--train-iters") + 1] = "4"
    plain = _losses(_run("pretrain_gpt.py", base))
    rec = _losses(_run("pretrain_gpt.py", base + ["--recompute-granularity", "full"]))
    assert sorted(plain) == [1, 2, 3, 4] and plain == rec
    # the reference's recompute methods: uniform groups of 2 layers, and block (first layer only)
    l4 = [a for a in base]
    l4[l4.index("--num-layers") + 1] = "4"
    plain4 = _losses(_run("pretrain_gpt.py", l4))
    for extra in (["--recompute-method", "uniform", "--recompute-num-layers", "2"], ["--recompute-method", "block", "--recompute-num-layers", "1"]):
        assert _losses(_run("pretrain_gpt.py", l4 + ["--recompute-granularity", "full"] + extra)) == plain4, extra


def test_gpt_distributed_saved_activations_tp2():
    """--distribute-saved-activations: the checkpointed layer input is kept as 1/t slices and re-gathered in backward"""
    base = [a for a in GPT]
    base[base.index("--train-iters") + 1] = "3"
    tp = ["--tensor-model-parallel-size", "2", "--hidden-dropout", "0", "--attention-dropout", "0"]
    plain = _losses(_run("pretrain_gpt.py", base + tp, nproc=2))
    rec = _losses(_run("pretrain_gpt.py", base + tp + ["--recompute-granularity", "full", "--recompute-method", "uniform",
                                                        "--distribute-saved-activations"], nproc=2))
    # two separate 2-rank runs: multi-threaded CPU GEMMs are not bit-reproducible across processes, so compare to rounding
    assert sorted(plain) == [1, 2, 3] == sorted(rec) and all(abs(float(plain[i]) - float(rec[i])) < 2e-5 for i in plain), (plain, rec)


def test_gpt_distributed_optimizer_matches_replicated_and_resumes(tmp_path):
    """--use-distributed-optimizer (optimizer state sharded over DP=2): same losses as the replicated optimizer, one
    optimizer shard file per DP rank, bit-exact resume."""
    base = [a for a in GPT]
    base[base.index("--global-batch-size") + 1] = "8"
    base += ["--hidden-dropout", "0", "--attention-dropout", "0"]
    plain = _losses(_run("pretrain_gpt.py", base, nproc=2))
    ck = str(tmp_path / "ck")
    dist_run = _losses(_run("pretrain_gpt.py", base + ["--use-distributed-optimizer", "--save", ck, "--save-interval", "3"], nproc=2))
    assert plain == dist_run and sorted(plain) == [1, 2, 3, 4, 5, 6]
    d = os.path.join(ck, "iter_0000003", "mp_rank_00")
    assert sorted(os.listdir(d)) == ["distrib_optim_dp000.pt", "distrib_optim_dp001.pt", "model_optim_rng.pt"]
    with open(os.path.join(ck, "latest_checkpointed_iteration.txt"), "w") as f:
        f.write("3")
    resumed = _losses(_run("pretrain_gpt.py", base + ["--use-distributed-optimizer", "--load", ck], nproc=2))
    assert sorted(resumed) == [4, 5, 6] and all(_close(resumed[i], plain[i]) for i in (4, 5, 6))


def test_gpt_pretrain_llama_style_flags_and_switch_mlp_resume(tmp_path):
    """--swiglu / RoPE / GQA / --disable-bias-linear / --num-experts go through the modular layer; checkpoints resume bit-exactly"""
    ck = str(tmp_path / "ck")
    extra = ["--swiglu", "--use-rotary-position-embeddings", "--group-query-attention", "--num-query-groups", "2",
             "--disable-bias-linear", "--untie-embeddings-and-output-weights", "--num-experts", "2"]
    out = _run("pretrain_gpt.py", GPT + extra + ["--save", ck, "--save-interval", "3"])
    full = _losses(out)
    assert sorted(full) == [1, 2, 3, 4, 5, 6]
    assert float(full[6]) < float(full[1])
    with open(os.path.join(ck, "latest_checkpointed_iteration.txt"), "w") as f:
        f.write("3")
    resumed = _losses(_run("pretrain_gpt.py", GPT + extra + ["--load", ck]))
    assert sorted(resumed) == [4, 5, 6] and all(_close(resumed[i], full[i]) for i in (4, 5, 6))


def test_gpt_pretrain_structured_metrics_and_fault_injection_recovery(tmp_path):
    """JSONL + scalar sinks, --log-params-norm / --log-num-zeros-in-grad, and the recovery story
    (a rank dies after a checkpoint -> relaunch with --load resumes from the latest checkpoint with identical losses)"""
    import json
    ck, mj, tb = str(tmp_path / "ck"), str(tmp_path / "m" / "metrics.jsonl"), str(tmp_path / "tb")
    obs = ["--metrics-jsonl", mj, "--tensorboard-dir", tb, "--log-params-norm", "--log-num-zeros-in-grad", "--timing-log-level", "1"]
    full = _losses(_run("pretrain_gpt.py", GPT + obs))
    recs = [json.loads(l) for l in open(mj)]
    assert [r["iteration"] for r in recs] == [1, 2, 3, 4, 5, 6]
    for r in recs:
        assert r["tokens_per_s"] > 0 and 0 < r["mfu"] < 1 and r["params_norm"] > 0 and r["num_zeros_in_grad"] >= 0 and "lm_loss" in r
        assert abs(r["lm_loss"] - float(full[r["iteration"]])) < 1e-5
    assert os.listdir(tb)
    # a fault after iteration 4 (checkpoint at 3): the job fails ...
    cmd = [sys.executable, os.path.join(ROOT, "entrypoints", "pretrain_gpt.py"), *GPT, "--save", ck, "--save-interval", "3", "--inject-fault", "0:4"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode != 0 and "injected fault on rank 0 at iteration 4" in r.stderr
    assert open(os.path.join(ck, "latest_checkpointed_iteration.txt")).read().strip() == "3"
    # ... and the relaunch picks up at iteration 4 with the losses of the uninterrupted run
    resumed = _losses(_run("pretrain_gpt.py", GPT + ["--load", ck, "--save", ck, "--save-interval", "3"]))
    assert sorted(resumed) == [4, 5, 6] and all(_close(resumed[i], full[i]) for i in (4, 5, 6))


def test_gpt_sample_based_training_with_batch_size_rampup():
    """--train-samples + --rampup-batch-size 2 2 8: global batch 2 -> 4 over the first 8 samples (megatron/microbatches.py)"""
    argv = [a for a in GPT]
    for flag in ("--train-iters", "--lr-decay-iters", "--lr-warmup-iters"):
        i = argv.index(flag)
        del argv[i:i + 2]
    argv[argv.index("--micro-batch-size") + 1] = "1"
    out = _run("pretrain_gpt.py", argv + ["--train-samples", "20", "--lr-decay-samples", "20", "--lr-warmup-samples", "4", "--rampup-batch-size", "2", "2", "8"])
    rows = re.findall(r"iteration\s+(\d+)/\s*(\d+) \| consumed samples:\s+(\d+) .*? global batch size:\s+(\d+) ", out)
    assert rows, out[-2000:]
    sizes = [int(r[3]) for r in rows]
    consumed = [int(r[2]) for r in rows]
    assert sizes[0] == 2 and sizes[-1] == 4 and sorted(sizes) == sizes
    assert all(consumed[i] - (consumed[i - 1] if i else 0) == sizes[i] for i in range(len(rows)))
    assert int(rows[0][1]) == len(rows) and consumed[-1] <= 20


def test_checkpoint_util_reshards_tp_and_pp(tmp_path):
    """tools/checkpoint_util.py: a TP=2 checkpoint (SwiGLU + GQA, so the packed layouts are exercised) evaluated under TP=2, then
    re-sharded to TP=1 and to TP=1 x PP=2 - the validation loss of the re-loaded model is the same in every layout"""
    def val_loss(out):
        m = re.search(r"validation loss at the end of training for val data \| lm loss value: ([0-9.E+-]+)", out)
        assert m, out[-3000:]
        return float(m.group(1))

    ck = str(tmp_path / "tp2")
    argv = [a for a in GPT]
    argv[argv.index("--train-iters") + 1] = "2"
    argv[argv.index("--num-layers") + 1] = "4"
    argv += ["--swiglu", "--group-query-attention", "--num-query-groups", "2", "--hidden-dropout", "0", "--attention-dropout", "0"]
    _run("pretrain_gpt.py", argv + ["--tensor-model-parallel-size", "2", "--save", ck, "--save-interval", "2"], nproc=2)
    ev = ["--finetune", "--skip-train"]
    ref = val_loss(_run("pretrain_gpt.py", argv + ["--tensor-model-parallel-size", "2", "--load", ck] + ev, nproc=2))
    tool = os.path.join(ROOT, "tools", "checkpoint_util.py")
    for tp, pp in ((1, 1), (1, 2)):
        out_dir = str(tmp_path / f"tp{tp}pp{pp}")
        r = subprocess.run([sys.executable, tool, "--load-dir", ck, "--save-dir", out_dir, "--target-tensor-parallel-size", str(tp),
                            "--target-pipeline-parallel-size", str(pp)], capture_output=True, text=True, timeout=300)
        assert r.returncode == 0, r.stdout + r.stderr
        extra = ["--tensor-model-parallel-size", str(tp), "--pipeline-model-parallel-size", str(pp), "--load", out_dir] + ev
        got = val_loss(_run("pretrain_gpt.py", argv + extra, nproc=tp * pp if tp * pp > 1 else 0))
        assert abs(got - ref) < 2e-4 * max(1.0, abs(ref)), (tp, pp, got, ref)
    # and back up: TP=1 -> TP=2 x PP=2 file layout
    out_dir = str(tmp_path / "tp2pp2")
    r = subprocess.run([sys.executable, tool, "--load-dir", str(tmp_path / "tp1pp1"), "--save-dir", out_dir, "--target-tensor-parallel-size", "2",
                        "--target-pipeline-parallel-size", "2"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert sorted(os.listdir(os.path.join(out_dir, "iter_0000002"))) == ["mp_rank_00_000", "mp_rank_00_001", "mp_rank_01_000", "mp_rank_01_001"]


def test_oxford_resnet_script_with_native_sync_batchnorm(tmp_path):
    """the script's default model family (--model_name resnet*, pytorch_oxford_ddp.py:123) with the head really replaced
    and every BatchNorm2d on the native SyncBatchNorm path"""
    out = _run("pytorch_oxford_ddp.py", ["--model_name", "resnet18", "--num-classes", "37", "--num-epochs", "1", "--batch-size", "4",
                                         "--test-batch-size", "2", "--synthetic", "--synthetic-train-size", "16", "--synthetic-val-size", "4",
                                         "--height", "64", "--width", "64", "--backend", "gloo", "--no-pretrained", "--num-workers", "0",
                                         "--model-dir", str(tmp_path), "--print-freq", "1", "--max-steps", "2", "--native-bn", "--sync-bn"], nproc=2, timeout=900)
    assert "Test_Prec@1=" in out and "Test_Loss=" in out
    import torch
    ck = torch.load(os.path.join(tmp_path, "checkpoint.pth"), map_location="cpu", weights_only=False)
    keys = list(ck["state_dict"])
    assert any(k.endswith("bn1.running_mean") for k in keys) and any(k.endswith("fc.weight") and ck["state_dict"][k].shape[0] == 37 for k in keys)


def test_elastic_restart_resumes_from_latest_checkpoint(tmp_path):
    """the recovery story of SURVEY 5.3 end to end: `torchrun --max-restarts 1`; rank 0 dies after iteration 4 of the first attempt,
    the agent restarts the workers, `--load` picks up the iteration-3 checkpoint and the job finishes with the uninterrupted losses"""
    ck = str(tmp_path / "ck")
    full = _losses(_run("pretrain_gpt.py", GPT, nproc=2))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--standalone", "--local-addr", "127.0.0.1", "--nnodes=1", "--nproc-per-node=2", "--max-restarts=1",
           os.path.join(ROOT, "entrypoints", "pretrain_gpt.py"), *GPT, "--save", ck, "--save-interval", "3",
           "--load", ck, "--inject-fault", "0:4"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900, env={**os.environ, "OMP_NUM_THREADS": "2", "MKL_CBWR": "AVX2"})
    out = r.stdout + r.stderr
    assert r.returncode == 0, out[-4000:]
    assert "injected fault on rank 0 at iteration 4" in out
    its = [int(m.group(1)) for m in re.finditer(r"iteration\s+(\d+)/", out)]
    assert its[:4] == [1, 2, 3, 4] and its[-3:] == [4, 5, 6]                 # second attempt starts after the iteration-3 checkpoint
    final = {int(m.group(1)): m.group(2) for m in re.finditer(r"iteration\s+(\d+)/\s*\d+ \|.*?lm loss: ([0-9.E+-]+)", out)}
    assert all(abs(float(final[i]) - float(full[i])) < 2e-5 for i in (4, 5, 6)), (final, full)          # same 2-rank job, uninterrupted


def test_gpt_fp16_dynamic_loss_scale_skips_then_trains():
    """--fp16 true (the notebook's spelling, Q10) with a too-large initial loss scale: overflowing iterations are skipped (no lm loss
    in the line, scheduler not advanced, scale halved), then training proceeds at a scale that fits - the shape of 3_...ipynb:4347-4348"""
    argv = [a for a in GPT]
    argv[argv.index("--train-iters") + 1] = "10"
    argv[argv.index("--lr-decay-iters") + 1] = "10"
    out = _run("pretrain_gpt.py", argv + ["--fp16", "true", "--initial-loss-scale", "16777216", "--hysteresis", "1"])
    rows = []
    for line in out.splitlines():
        m = re.search(r"iteration\s+(\d+)/.*?learning rate: ([0-9.E+-]+) .*?loss scale: ([0-9.]+) .*?number of skipped iterations:\s+(\d+)", line)
        if m:
            rows.append((int(m.group(1)), float(m.group(2)), float(m.group(3)), int(m.group(4)), "lm loss:" in line))
    assert len(rows) == 10
    assert rows[0][3] == 1 and not rows[0][4] and rows[0][1] == 0.0 and rows[0][2] == 8388608.0        # skipped: no loss, lr not advanced, scale halved
    assert rows[-1][3] == 0 and rows[-1][4] and rows[-1][2] < 8388608.0 and rows[-1][1] > 0.0
    scales = [r[2] for r in rows]
    assert scales == sorted(scales, reverse=True)
    assert sum(r[3] for r in rows) >= 3


def test_gpt_tensorboard_switches(tmp_path):
    """--tensorboard-dir with the reference's scalar switches (arguments.py:652-679): timers (a collective over both pipeline ranks),
    batch size, world size, validation perplexity on; loss scale off"""
    tb = str(tmp_path / "tb")
    _run("pretrain_gpt.py", GPT + ["--tensorboard-dir", tb, "--tensorboard-log-interval", "1", "--log-timers-to-tensorboard", "--log-batch-size-to-tensorboard",
                                   "--log-world-size-to-tensorboard", "--log-validation-ppl-to-tensorboard", "--no-log-loss-scale-to-tensorboard",
                                   "--timing-log-level", "2", "--pipeline-model-parallel-size", "2", "--eval-interval", "3"], nproc=2)
    tags = set()
    if os.path.exists(os.path.join(tb, "scalars.jsonl")):          # no tensorboard package: the JSONL fallback of utils.metrics.ScalarWriter
        import json
        tags = {json.loads(l)["tag"] for l in open(os.path.join(tb, "scalars.jsonl"))}
    else:
        from tensorboard.backend.event_processing.event_accumulator import EventAccumulator
        ea = EventAccumulator(tb)
        ea.Reload()
        tags = set(ea.Tags()["scalars"])
    for want in ("lm-loss", "learning-rate", "global-batch-size", "world-size", "forward-backward-time", "optimizer-time", "lm loss validation",
                 "lm loss validation ppl"):
        assert want in tags, (want, sorted(tags))
    assert "loss-scale" not in tags
