"""2_training_oxford-pet_ddp.ipynb as a script (cells :426-435 metric_definitions, :446-458 hyper-parameters, :593 fit).
Data: `python tools/prepare_oxford_pet.py --images <oxford-iiit-pet/images> --output <dir>` then `--data <dir>`."""
from _common import ROOT, nproc, on_gpu, parse
import os

a = parse(__doc__)
from b200dist.launch.estimator import PyTorch

metric_definitions = [{"Name": "train:Time", "Regex": "Train_Time=(.*?):"}, {"Name": "train:Loss", "Regex": "Train_Loss=(.*?):"},
                      {"Name": "train:Prec@1", "Regex": "Train_Prec@1=(.*?):"}, {"Name": "train:Prec@5", "Regex": "Train_Prec@5=(.*?):"},
                      {"Name": "test:Time", "Regex": "Test_Time=(.*?):"}, {"Name": "test:Loss", "Regex": "Test_Loss=(.*?):"},
                      {"Name": "test:Prec@1", "Regex": "Test_Prec@1=(.*?):"}, {"Name": "test:Prec@5", "Regex": "Test_Prec@5=(.*?):"}]
if a.full:
    hyperparameters = {"model_name": "swin_b", "num-classes": 37, "height": 128, "width": 128, "num-epochs": 15, "batch-size": 80, "lr": 1e-4,
                       "backend": "smddp", "bf16": True}
    if not a.data:
        hyperparameters["synthetic"] = None
else:
    hyperparameters = {"model_name": "swin_t", "num-classes": 37, "height": 64, "width": 64, "num-epochs": 1, "batch-size": 2, "test-batch-size": 2,
                       "backend": "smddp" if on_gpu() else "gloo", "synthetic": None, "synthetic-train-size": 16, "synthetic-val-size": 4,
                       "no-pretrained": None, "num-workers": 0, "max-steps": 4, "print-freq": 1}
estimator = PyTorch(entry_point="pytorch_oxford_ddp.py", source_dir=os.path.join(ROOT, "entrypoints"), role=None, instance_count=1,
                    instance_type="local_gpu", processes_per_host=nproc(a), hyperparameters=hyperparameters,
                    distribution={"smdistributed": {"dataparallel": {"enabled": True}}}, metric_definitions=metric_definitions,
                    output_path=a.output, base_job_name="oxford-pet-ddp")
metrics = estimator.fit({"training": a.data} if a.data else None)
print({k: v[-1] for k, v in metrics.items() if v}, "model dir:", estimator.model_data)
