"""The notebooks' launch contract on one machine: `PyTorch(...).fit()` -> hyper-parameters as argv,
SM_* / OMPI_* environment, merged log, metric regexes, fail-stop."""
import json
import os
import sys
import textwrap

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_hyperparameters_become_string_argv():
    from b200dist.launch.estimator import hyperparameters_to_argv
    assert hyperparameters_to_argv({"epochs": 5, "backend": "nccl", "bf16": True, "lr": 1e-4}) == \
        ["--epochs", "5", "--backend", "nccl", "--bf16", "True", "--lr", "0.0001"]


def test_mnist_notebook_cell_runs_locally_with_mpi_style_env(tmp_path):
    """1_training_mnist_ddp.ipynb:302-306,329-351,440-455: same constructor arguments, two processes, gloo on CPU"""
    from b200dist.launch.estimator import PyTorch
    data = tmp_path / "data"
    data.mkdir()
    est = PyTorch(entry_point="pytorch_mnist_ddp.py", source_dir=os.path.join(ROOT, "entrypoints"), role="ignored", framework_version="1.13.1",
                  py_version="py39", instance_count=1, instance_type="local_gpu", processes_per_host=2, disable_profiler=True,
                  hyperparameters={"epochs": 1, "backend": "gloo", "synthetic": None, "synthetic-size": 256, "max-steps": 3, "log-interval": 1},
                  distribution={"mpi": {"enabled": True}}, output_path=str(tmp_path / "jobs"),
                  metric_definitions=[{"Name": "train:Loss", "Regex": r"Loss: ([0-9.]+)"}, {"Name": "test:Loss", "Regex": "Average loss: (.*?),"},
                                      {"Name": "test:Accuracy", "Regex": r"\((\d+)%\)"}])
    assert est.launcher_kind() == "mpi"
    assert est.command()[2:] == ["--epochs", "1", "--backend", "gloo", "--synthetic", "--synthetic-size", "256", "--max-steps", "3", "--log-interval", "1"]
    metrics = est.fit({"training": str(data)}, logs=False)
    assert len(metrics["train:Loss"]) >= 2 and len(metrics["test:Loss"]) == 1 and 0 <= metrics["test:Accuracy"][0] <= 100
    assert os.path.isdir(est.model_data) and "[mpi] 2 process(es)" in est.logs()


def test_environment_contract_and_fail_stop(tmp_path):
    from b200dist.launch.estimator import LocalEstimator, TrainingJobError
    script = tmp_path / "probe.py"
    script.write_text(textwrap.dedent('''
        import json, os, sys, time
        keys = ["SM_MODEL_DIR", "SM_CHANNEL_TRAINING", "SM_NUM_GPUS", "SM_HPS", "RANK", "WORLD_SIZE", "LOCAL_RANK", "MASTER_ADDR", "OMPI_COMM_WORLD_RANK"]
        print("ENV " + json.dumps({k: os.environ.get(k) for k in keys}) + " ARGV " + json.dumps(sys.argv[1:]), flush=True)
        if "--fail" in sys.argv and os.environ["RANK"] == "1":
            sys.exit(3)
        if "--fail" in sys.argv:
            time.sleep(60)
        print("value=%d;" % (10 + int(os.environ["RANK"])))
    '''))
    ch = tmp_path / "ch"
    ch.mkdir()
    est = LocalEstimator(entry_point=str(script), hyperparameters={"alpha": 0.5, "flag": True}, processes_per_host=2, output_path=str(tmp_path / "j"),
                         metric_definitions=[{"Name": "v", "Regex": "value=(.*?);"}], distribution={"torch_distributed": {"enabled": True}})
    m = est.fit(str(ch), logs=False)
    assert sorted(m["v"]) == [10.0, 11.0]
    envs = [json.loads(l.split("ENV ")[1].split(" ARGV ")[0]) for l in est.logs().splitlines() if l.startswith("ENV ")]
    assert sorted(e["RANK"] for e in envs) == ["0", "1"] and all(e["WORLD_SIZE"] == "2" and e["MASTER_ADDR"] == "127.0.0.1" for e in envs)
    assert all(e["SM_CHANNEL_TRAINING"] == str(ch) and e["SM_NUM_GPUS"] == "2" and e["OMPI_COMM_WORLD_RANK"] is None for e in envs)
    assert json.loads(envs[0]["SM_HPS"]) == {"alpha": 0.5, "flag": True}
    assert '["--alpha", "0.5", "--flag", "True"]' in est.logs()
    bad = LocalEstimator(entry_point=str(script), hyperparameters={"fail": ""}, processes_per_host=2, output_path=str(tmp_path / "j2"))
    bad.command = lambda: [sys.executable, str(script), "--fail"]
    import time
    t0 = time.time()
    with pytest.raises(TrainingJobError, match="rank 1 exited with code 3"):
        bad.fit(logs=False)
    assert time.time() - t0 < 30          # the sleeping rank was taken down, not waited for
    with pytest.raises(ValueError):
        LocalEstimator(entry_point=str(script), instance_count=2)


def test_example_script_of_notebook_1_runs(tmp_path):
    import subprocess
    r = subprocess.run([sys.executable, os.path.join(ROOT, "examples", "1_mnist_ddp.py"), "--processes", "2", "--output", str(tmp_path)],
                       capture_output=True, text=True, timeout=600, cwd=os.path.join(ROOT, "examples"))
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "'test:Accuracy'" in r.stdout and "model:" in r.stdout
