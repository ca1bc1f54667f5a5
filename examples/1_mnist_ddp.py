"""1_training_mnist_ddp.ipynb as a script (cells :302-306 metric_definitions, :329-351 hyper-parameters / distribution, :440-492 fit)."""
from _common import ROOT, nproc, on_gpu, parse
import os

a = parse(__doc__)
from b200dist.launch.estimator import PyTorch

metric_definitions = [{"Name": "train:Loss", "Regex": r"Loss: ([0-9.]+)"}, {"Name": "test:Loss", "Regex": "Average loss: (.*?),"},
                      {"Name": "test:Accuracy", "Regex": r"\((\d+)%\)"}]
hyperparameters = {"epochs": 5, "backend": "nccl"} if a.full else \
    {"epochs": 1, "backend": "nccl" if on_gpu() else "gloo", "synthetic": None, "synthetic-size": 2048, "log-interval": 4}
distribution = {"mpi": {"enabled": True}}          # the notebook's choice for backend == 'nccl'; {"smdistributed": ...} for 'smddp'
estimator = PyTorch(entry_point="pytorch_mnist_ddp.py", source_dir=os.path.join(ROOT, "entrypoints"), role=None, framework_version="1.13.1",
                    py_version="py39", instance_count=1, instance_type="local_gpu", processes_per_host=nproc(a), hyperparameters=hyperparameters,
                    distribution=distribution, metric_definitions=metric_definitions, disable_profiler=True, debugger_hook_config=False,
                    output_path=a.output, base_job_name="mnist-ddp")
inputs = {"training": a.data} if a.data else None
metrics = estimator.fit(inputs)
print({k: v[-1] for k, v in metrics.items() if v}, "model:", estimator.model_data)
