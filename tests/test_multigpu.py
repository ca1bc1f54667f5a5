"""Multi-GPU tests (need >= 2 B200s on one box; skipped otherwise): the fabric collectives, the fused AG->GEMM / GEMM->RS
kernels and the ZeRO-3 engine on the native path, each launched with torchrun inside a timeout and compared against NCCL."""
import os
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.gpu


def _ngpu():
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


def _torchrun(script, nproc, timeout, *argv):
    cmd = ["timeout", "-k", "5", str(timeout), sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
           "--master-addr", "127.0.0.1", "--master-port", str(29700 + os.getpid() % 200), os.path.join(ROOT, script), *argv]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout + 30)
    return r.returncode, r.stdout + r.stderr


@pytest.mark.skipif(_ngpu() < 2, reason="needs 2 GPUs")
def test_fabric_collectives_and_fused_gemm_kernels():
    rc, out = _torchrun("bench/comm_check.py", 2, 240)
    assert rc == 0 and "ALL OK" in out and "FAIL " not in out, out[-4000:]


@pytest.mark.skipif(_ngpu() < 2, reason="needs 2 GPUs")
def test_zero3_native_fabric_matches_nccl():
    rc, out = _torchrun("tests/_mgpu_zero3.py", 2, 300)
    assert rc == 0 and "ALL OK" in out, out[-4000:]


@pytest.mark.skipif(_ngpu() < 2, reason="needs 2 GPUs")
def test_gpt_tp2_sp_native_matches_nccl_backend():
    """same seed, same data: the fused-kernel TP path and the NCCL TP path produce the same loss curve"""
    base = ["--num-layers", "2", "--hidden-size", "256", "--num-attention-heads", "4", "--seq-length", "256", "--max-position-embeddings", "256",
            "--micro-batch-size", "4", "--global-batch-size", "8", "--lr", "1e-3", "--train-iters", "4", "--lr-decay-iters", "4", "--lr-warmup-iters", "1",
            "--log-interval", "1", "--eval-interval", "100", "--eval-iters", "0", "--tokenizer-type", "NullTokenizer", "--vocab-size", "1023", "--mock-data",
            "--bf16", "--tensor-model-parallel-size", "2", "--sequence-parallel", "--seed", "11", "--hidden-dropout", "0", "--attention-dropout", "0"]
    import re
    curves = {}
    for backend in ("native", "nccl"):
        rc, out = _torchrun("entrypoints/pretrain_gpt.py", 2, 300, *base, "--comm-backend", backend)
        assert rc == 0, out[-4000:]
        curves[backend] = [float(x) for x in re.findall(r"lm loss: ([0-9.E+-]+)", out)]
    assert len(curves["native"]) == 4 == len(curves["nccl"])
    assert max(abs(a - b) for a, b in zip(curves["native"], curves["nccl"])) < 5e-2, curves


@pytest.mark.skipif(_ngpu() < 2, reason="needs 2 GPUs")
def test_gpt_tp2_without_sp_gemm_allreduce_kernel_matches_nccl_backend():
    """TP=2 WITHOUT sequence parallelism (the reference's logged layout): RowParallelLinear forward / ColumnParallelLinear
    backward run the fused GEMM->all-reduce kernel; same loss curve as GEMM + NCCL all_reduce"""
    base = ["--num-layers", "2", "--hidden-size", "256", "--num-attention-heads", "4", "--seq-length", "256", "--max-position-embeddings", "256",
            "--micro-batch-size", "4", "--global-batch-size", "8", "--lr", "1e-3", "--train-iters", "4", "--lr-decay-iters", "4", "--lr-warmup-iters", "1",
            "--log-interval", "1", "--eval-interval", "100", "--eval-iters", "0", "--tokenizer-type", "NullTokenizer", "--vocab-size", "1023", "--mock-data",
            "--bf16", "--tensor-model-parallel-size", "2", "--seed", "11", "--hidden-dropout", "0", "--attention-dropout", "0"]
    import re
    curves = {}
    for backend in ("native", "nccl"):
        rc, out = _torchrun("entrypoints/pretrain_gpt.py", 2, 300, *base, "--comm-backend", backend)
        assert rc == 0, out[-4000:]
        curves[backend] = [float(x) for x in re.findall(r"lm loss: ([0-9.E+-]+)", out)]
    assert len(curves["native"]) == 4 == len(curves["nccl"])
    assert max(abs(a - b) for a, b in zip(curves["native"], curves["nccl"])) < 5e-2, curves


@pytest.mark.skipif(_ngpu() < 2, reason="needs 2 GPUs")
def test_gpt_pp2_batched_p2p_matches_single_gpu():
    """PP=2 (1F1B, then interleaved) over NCCL with the batched p2p transport against the same model on one GPU; every
    launch sits inside `timeout`, and B2_HANG_DUMP_S prints where each rank waits if it stalls"""
    import re
    base = ["--num-layers", "4", "--hidden-size", "256", "--num-attention-heads", "4", "--seq-length", "256", "--max-position-embeddings", "256",
            "--micro-batch-size", "2", "--global-batch-size", "8", "--lr", "1e-3", "--train-iters", "3", "--lr-decay-iters", "3", "--lr-warmup-iters", "1",
            "--log-interval", "1", "--eval-interval", "100", "--eval-iters", "0", "--tokenizer-type", "NullTokenizer", "--vocab-size", "1023", "--mock-data",
            "--bf16", "--seed", "11", "--hidden-dropout", "0", "--attention-dropout", "0"]
    os.environ["B2_HANG_DUMP_S"] = "60"
    try:
        rc, out = _torchrun("entrypoints/pretrain_gpt.py", 1, 240, *base)
        assert rc == 0, out[-4000:]
        ref = [float(x) for x in re.findall(r"lm loss: ([0-9.E+-]+)", out)]
        for extra in ([], ["--num-layers-per-virtual-pipeline-stage", "1"]):
            rc, out = _torchrun("entrypoints/pretrain_gpt.py", 2, 240, *base, "--pipeline-model-parallel-size", "2", *extra)
            assert rc == 0, out[-6000:]
            got = [float(x) for x in re.findall(r"lm loss: ([0-9.E+-]+)", out)]
            assert len(got) == 3 == len(ref) and max(abs(a - b) for a, b in zip(got, ref)) < 5e-2, (extra, got, ref)
    finally:
        os.environ.pop("B2_HANG_DUMP_S", None)


@pytest.mark.skipif(_ngpu() < 4, reason="needs 4 GPUs")
def test_gpt_tp2_pp2_interleaved_fused_tp_kernels_match_nccl_tp():
    """TP=2 (+SP) x PP=2 interleaved on 4 GPUs: the fused AG->GEMM / GEMM->RS kernels next to NCCL pipeline p2p
    (B2_PP_NATIVE_TP=1: SMs reserved for the NCCL send/recv kernels) give the loss curve of the all-NCCL run"""
    import re
    base = ["--num-layers", "4", "--hidden-size", "256", "--num-attention-heads", "4", "--seq-length", "256", "--max-position-embeddings", "256",
            "--micro-batch-size", "2", "--global-batch-size", "8", "--lr", "1e-3", "--train-iters", "3", "--lr-decay-iters", "3", "--lr-warmup-iters", "1",
            "--log-interval", "1", "--eval-interval", "100", "--eval-iters", "0", "--tokenizer-type", "NullTokenizer", "--vocab-size", "1023", "--mock-data",
            "--bf16", "--seed", "11", "--hidden-dropout", "0", "--attention-dropout", "0", "--tensor-model-parallel-size", "2", "--sequence-parallel",
            "--pipeline-model-parallel-size", "2", "--num-layers-per-virtual-pipeline-stage", "1"]
    curves = {}
    os.environ["B2_HANG_DUMP_S"] = "60"
    try:
        for native in ("0", "1"):
            os.environ["B2_PP_NATIVE_TP"] = native
            rc, out = _torchrun("entrypoints/pretrain_gpt.py", 4, 240, *base, "--comm-backend", "native")
            assert rc == 0, out[-6000:]
            if native == "1":
                assert "reserving SMs for NCCL p2p" in out, out[-3000:]
            curves[native] = [float(x) for x in re.findall(r"lm loss: ([0-9.E+-]+)", out)]
    finally:
        os.environ.pop("B2_PP_NATIVE_TP", None)
        os.environ.pop("B2_HANG_DUMP_S", None)
    assert len(curves["0"]) == 3 == len(curves["1"])
    assert max(abs(a - b) for a, b in zip(curves["0"], curves["1"])) < 5e-2, curves


@pytest.mark.skipif(_ngpu() < 2, reason="needs 2 GPUs")
def test_distributed_optimizer_fabric_collectives_match_nccl():
    """--use-distributed-optimizer on DP=2: gradient reduce-scatter and parameter all-gather on the fabric kernels (symmetric
    main_grad / model arenas) against reduce_scatter_tensor / all_gather_into_tensor"""
    import re
    base = ["--num-layers", "2", "--hidden-size", "256", "--num-attention-heads", "4", "--seq-length", "256", "--max-position-embeddings", "256",
            "--micro-batch-size", "4", "--global-batch-size", "16", "--lr", "1e-3", "--train-iters", "4", "--lr-decay-iters", "4", "--lr-warmup-iters", "1",
            "--log-interval", "1", "--eval-interval", "100", "--eval-iters", "0", "--tokenizer-type", "NullTokenizer", "--vocab-size", "1023", "--mock-data",
            "--bf16", "--seed", "11", "--hidden-dropout", "0", "--attention-dropout", "0", "--use-distributed-optimizer"]
    curves = {}
    for backend in ("native", "nccl"):
        rc, out = _torchrun("entrypoints/pretrain_gpt.py", 2, 300, *base, "--comm-backend", backend)
        assert rc == 0, out[-4000:]
        curves[backend] = [float(x) for x in re.findall(r"lm loss: ([0-9.E+-]+)", out)]
    assert len(curves["native"]) == 4 == len(curves["nccl"])
    assert max(abs(a - b) for a, b in zip(curves["native"], curves["nccl"])) < 5e-2, curves
