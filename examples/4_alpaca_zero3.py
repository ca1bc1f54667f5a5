"""4_training_alpaca_deepspeed.ipynb as a script (cells :464-485 hyper-parameters, :600-616 estimator, :663 fit).
--full: LLaMA-7B shape with the reference's DeepSpeed JSON (`--data alpaca_data.json`; without it: synthetic instruction batches)."""
from _common import ROOT, nproc, on_gpu, parse
import os

a = parse(__doc__)
from b200dist.launch.estimator import PyTorch

cfg = os.path.join(ROOT, "configs", "default_offload_opt_param.json")
if a.full:
    hyperparameters = {"model_name_or_path": "llama-7b", "bf16": True, "output_dir": os.path.join(a.output, "alpaca_out"), "num_train_epochs": 1,
                       "per_device_train_batch_size": 4, "per_device_eval_batch_size": 4, "gradient_accumulation_steps": 8, "evaluation_strategy": "no",
                       "save_strategy": "steps", "save_steps": 2000, "save_total_limit": 1, "learning_rate": 2e-5, "weight_decay": 0., "warmup_ratio": 0.03,
                       "deepspeed": cfg, "tf32": True, "report_to": "none"}
    hyperparameters.update({"data_path": a.data} if a.data else {"synthetic_data": True})
else:
    hyperparameters = {"model_name_or_path": "opt-tiny", "bf16": on_gpu(), "output_dir": os.path.join(a.output, "alpaca_out"), "max_steps": 6,
                       "per_device_train_batch_size": 1, "gradient_accumulation_steps": 2, "learning_rate": 1e-3, "logging_steps": 1, "model_max_length": 64,
                       "deepspeed": cfg, "synthetic_data": True, "synthetic_samples": 64, "save_strategy": "no", "report_to": "none"}
metric_definitions = [{"Name": "loss", "Regex": r"'loss': ([0-9.]+)"}, {"Name": "train_samples_per_second", "Regex": r"'train_samples_per_second': ([0-9.]+)"}]
estimator = PyTorch(entry_point="train.py", source_dir=os.path.join(ROOT, "entrypoints"), role=None, instance_count=1, instance_type="local_gpu",
                    processes_per_host=nproc(a), hyperparameters=hyperparameters, distribution={"mpi": {"enabled": True}},
                    metric_definitions=metric_definitions, output_path=a.output, base_job_name="alpaca-zero3")
metrics = estimator.fit()
print({k: v[-1] for k, v in metrics.items() if v})
