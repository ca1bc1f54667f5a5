"""3_training_megatron-lm.ipynb as a script (cells :400-427 hyper-parameters, :439 mpi distribution, :607 fit).
--full: GPT-2 small, TP=4 as in the notebook (needs >= 4 GPUs and --data <prefix of the .bin/.idx files> with vocab.json / merges.txt next to it)."""
from _common import ROOT, nproc, on_gpu, parse
import os

a = parse(__doc__)
from b200dist.launch.estimator import PyTorch

if a.full:
    d = a.data or "/data/codeparrot_content_document"
    hyperparameters = {"num-layers": 12, "hidden-size": 768, "num-attention-heads": 12, "seq-length": 1024, "max-position-embeddings": 1024,
                       "micro-batch-size": 12, "global-batch-size": 192, "lr": 0.0005, "train-iters": 150000, "lr-decay-iters": 150000,
                       "lr-decay-style": "cosine", "lr-warmup-iters": 2000, "weight-decay": .1, "adam-beta2": .999, "fp16": "true", "log-interval": 10,
                       "save-interval": 2000, "eval-interval": 200, "eval-iters": 10, "data-path": d, "vocab-file": os.path.join(os.path.dirname(d), "vocab.json"),
                       "merge-file": os.path.join(os.path.dirname(d), "merges.txt"), "save": os.path.join(a.output, "gpt_ckpt"),
                       "load": os.path.join(a.output, "gpt_ckpt"), "tensor-model-parallel-size": 4, "pipeline-model-parallel-size": 1}
else:
    n = nproc(a)
    hyperparameters = {"num-layers": 2, "hidden-size": 64, "num-attention-heads": 4, "seq-length": 32, "max-position-embeddings": 32, "micro-batch-size": 2,
                       "global-batch-size": 4 * max(1, n // 2), "lr": 1e-3, "train-iters": 8, "lr-decay-iters": 8, "lr-decay-style": "cosine", "lr-warmup-iters": 2,
                       "weight-decay": .1, "log-interval": 1, "eval-interval": 100, "eval-iters": 1, "tokenizer-type": "NullTokenizer", "vocab-size": 255,
                       "mock-data": None, "tensor-model-parallel-size": 2 if n >= 2 else 1, "sequence-parallel": None,
                       "distributed-backend": "nccl" if on_gpu() else "gloo"}
metric_definitions = [{"Name": "lm loss", "Regex": r"lm loss: ([0-9.E+-]+)"}, {"Name": "tokens/s", "Regex": r"tokens/s: (\d+)"}]
estimator = PyTorch(entry_point="pretrain_gpt.py", source_dir=os.path.join(ROOT, "entrypoints"), role=None, instance_count=1, instance_type="local_gpu",
                    processes_per_host=nproc(a), hyperparameters=hyperparameters, distribution={"mpi": {"enabled": True}},
                    metric_definitions=metric_definitions, output_path=a.output, base_job_name="megatron-gpt")
metrics = estimator.fit()
print({k: v[-1] for k, v in metrics.items() if v})
