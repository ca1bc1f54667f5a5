"""HF GPT-2 <-> Megatron-layout conversion (tools/convert_gpt2.py): the flagship GPT model, fed the converted weights of a real
`transformers.GPT2LMHeadModel`, reproduces its logits and loss - an independent-implementation parity check of the hand-scheduled
layer (QKV per-head layout, tanh GeLU, LayerNorm, tied head) - and the entry point evaluates a converted checkpoint."""
import importlib.util
import os
import re
import subprocess
import sys

import pytest
import torch

transformers = pytest.importorskip("transformers")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _tool(name):
    spec = importlib.util.spec_from_file_location(name, os.path.join(ROOT, "tools", name + ".py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def _hf_gpt2():
    torch.manual_seed(0)
    cfg = transformers.GPT2Config(vocab_size=256, n_positions=32, n_embd=64, n_layer=2, n_head=4, resid_pdrop=0.0, embd_pdrop=0.0, attn_pdrop=0.0,
                                  activation_function="gelu_new")
    m = transformers.GPT2LMHeadModel(cfg).eval()
    with torch.no_grad():
        for p in m.parameters():
            if p.dim() == 1:
                p.add_(torch.randn_like(p) * 0.05)       # HF initialises biases to zero: make them count
    return m


def test_gpt_model_reproduces_hf_gpt2_with_converted_weights(tmp_path):
    from b200dist.config.megatron_args import TransformerConfig
    from b200dist.models.gpt import GPTModel
    conv = _tool("convert_gpt2")
    hf = _hf_gpt2()
    flat = conv.hf_gpt2_to_megatron(hf.state_dict(), 2, 4, 256)
    cfg = TransformerConfig(num_layers=2, hidden_size=64, num_attention_heads=4, ffn_hidden_size=256, kv_channels=16, seq_length=32,
                            max_position_embeddings=32, padded_vocab_size=256, hidden_dropout=0.0, attention_dropout=0.0, params_dtype=torch.float32,
                            comm_backend="nccl")
    ours = GPTModel(cfg, device="cpu", layer_offset=0, num_layers_here=2).eval()
    ours.load_state_dict(flat)
    ids = torch.randint(0, 256, (3, 32))
    with torch.no_grad():
        want = hf(input_ids=ids).logits                                  # [B, S, V]
        got = ours(ids).permute(1, 0, 2)                                  # ours: [S, B, V]
        assert torch.allclose(got, want, atol=2e-4, rtol=1e-3), (got - want).abs().max()
        labels = torch.randint(0, 256, (3, 32))
        ref_loss = torch.nn.functional.cross_entropy(want.reshape(-1, 256), labels.reshape(-1), reduction="none").view(3, 32)
        assert torch.allclose(ours(ids, labels=labels), ref_loss, atol=2e-4, rtol=1e-3)
    # and back: Megatron layout -> HF names, bit-identical tensors
    back = conv.megatron_to_hf_gpt2(ours.state_dict(), 2, 4, 256)
    for k, v in hf.state_dict().items():
        if k.endswith((".attn.bias", ".attn.masked_bias")):
            continue
        assert torch.equal(back[k], v), k


def test_converted_checkpoint_runs_through_the_entry_point_and_resharding(tmp_path):
    conv = _tool("convert_gpt2")
    hf = _hf_gpt2()
    hf_dir, ck = str(tmp_path / "hf"), str(tmp_path / "meg")
    hf.save_pretrained(hf_dir, safe_serialization=True)
    conv.convert_hf_dir_to_megatron(hf_dir, ck)
    assert open(os.path.join(ck, "latest_checkpointed_iteration.txt")).read() == "release"
    common = ("--num-layers 2 --hidden-size 64 --num-attention-heads 4 --seq-length 32 --max-position-embeddings 32 --micro-batch-size 2 "
              "--global-batch-size 4 --lr 1e-3 --log-interval 1 --eval-interval 100 --eval-iters 2 --tokenizer-type NullTokenizer --vocab-size 255 "
              "--mock-data --distributed-backend gloo --seed 7 --train-iters 2 --hidden-dropout 0 --attention-dropout 0 --finetune --skip-train").split()

    def val(cmd):
        r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env={**os.environ, "OMP_NUM_THREADS": "2"})
        assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]
        m = re.search(r"validation loss at the end of training for val data \| lm loss value: ([0-9.E+-]+)", r.stdout + r.stderr)
        assert m, r.stdout[-2000:]
        return float(m.group(1))

    entry = os.path.join(ROOT, "entrypoints", "pretrain_gpt.py")
    one = val([sys.executable, entry, *common, "--load", ck])
    assert 4.0 < one < 8.0                      # random GPT-2 on random tokens: about ln(256) = 5.5
    tp2 = str(tmp_path / "tp2")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "checkpoint_util.py"), "--load-dir", ck, "--save-dir", tp2,
                        "--target-tensor-parallel-size", "2"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    with __import__("socket").socket() as so:
        so.bind(("127.0.0.1", 0))
        port = so.getsockname()[1]
    two = val([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1", "--master-port", str(port),
               entry, *common, "--tensor-model-parallel-size", "2", "--load", tp2])
    assert abs(one - two) < 2e-4 * one
    # Megatron -> HF directory that transformers can open
    out_hf = str(tmp_path / "hf_out")
    conv.convert_megatron_dir_to_hf(tp2, out_hf, vocab_size=256)
    again = transformers.GPT2LMHeadModel.from_pretrained(out_hf).eval()
    ids = torch.randint(0, 256, (2, 16))
    with torch.no_grad():
        assert torch.allclose(again(input_ids=ids).logits, hf(input_ids=ids).logits, atol=1e-5)
