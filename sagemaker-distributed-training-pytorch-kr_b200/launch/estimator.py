"""Local stand-in for the launch path of the reference's notebooks: `PyTorch(entry_point, source_dir, hyperparameters,
distribution, metric_definitions, ...).fit(inputs)` (1_training_mnist_ddp.ipynb:440-455,492 and the three other notebooks;
SURVEY 3.1, T17-T20, E32).

On SageMaker that call provisions instances, downloads the channels and has the training toolkit start
`mpirun ... python <entry>.py --key value ...` with one process per GPU and the `SM_*` / `OMPI_COMM_WORLD_*` environment.
Here the same constructor arguments start the same entry point on THIS machine, one process per GPU, and reproduce the parts
of the contract that user code can observe:

* hyper-parameters become `--key value` argv, booleans as the strings `True` / `False` (4_...ipynb:877) - `SM_USER_ARGS`, `SM_HPS`;
* channels are local directories exported as `SM_CHANNEL_<NAME>` (+ `SM_CHANNELS`), the model directory as `SM_MODEL_DIR`,
  plus `SM_NUM_GPUS`, `SM_HOSTS`, `SM_CURRENT_HOST`, `SM_OUTPUT_DATA_DIR`;
* `distribution={"mpi": ...}` or `{"smdistributed": {"dataparallel": ...}}` (1_...ipynb:344-351): processes get
  `OMPI_COMM_WORLD_{SIZE,RANK,LOCAL_RANK}` like the toolkit's mpirun; `{"torch_distributed": {"enabled": True}}` or nothing:
  `RANK / LOCAL_RANK / WORLD_SIZE` like torchrun.  Rendezvous is always 127.0.0.1.  NCCL_DEBUG=INFO is exported as the
  toolkit does (1_...ipynb:686) unless the caller's environment sets it;
* `metric_definitions=[{"Name": ..., "Regex": ...}]` (1_...ipynb:302-306) are applied to the merged log; `fit()` returns
  the extracted series, what the SageMaker console would chart;
* fail-stop: the first rank that exits non-zero takes the job down (mpirun `orte_abort_on_non_zero_status`), `max_run`
  seconds is the wall-clock limit (1_...ipynb:365).

Instance types, roles, images and S3 are accepted and ignored; `instance_count` must be 1 (one box; multi-node is a
torchrun / mpirun command line, see README).
"""
from __future__ import annotations

import json
import os
import re
import shlex
import signal
import socket
import subprocess
import sys
import threading
import time
from typing import Dict, List, Optional, Union


def hyperparameters_to_argv(hp: Optional[Dict]) -> List[str]:
    """the training toolkit's mapping: {'epochs': 5, 'bf16': True} -> ['--epochs', '5', '--bf16', 'True'];
    (addition) a value of None emits the bare flag, for argparse `store_true` options"""
    argv: List[str] = []
    for k, v in (hp or {}).items():
        argv.append("--" + str(k))
        if v is None:
            continue
        if isinstance(v, (dict, list)):
            argv.append(json.dumps(v))
        else:
            argv.append(str(v))
    return argv


def _free_port() -> int:
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


class TrainingJobError(RuntimeError):
    pass


class LocalEstimator:
    def __init__(self, entry_point: str, source_dir: Optional[str] = None, hyperparameters: Optional[Dict] = None,
                 distribution: Optional[Dict] = None, metric_definitions: Optional[List[Dict]] = None, instance_count: int = 1,
                 instance_type: str = "local_gpu", processes_per_host: Optional[int] = None, max_run: Optional[int] = None,
                 model_dir: Optional[str] = None, output_path: Optional[str] = None, environment: Optional[Dict[str, str]] = None,
                 base_job_name: str = "b200dist", python: str = sys.executable, **ignored):
        if instance_count != 1:
            raise ValueError("LocalEstimator runs on one machine (instance_count=1); launch multi-node jobs with torchrun / mpirun")
        self.entry_point, self.source_dir = entry_point, source_dir
        self.hyperparameters = dict(hyperparameters or {})
        self.distribution = distribution or {}
        self.metric_definitions = list(metric_definitions or [])
        self.instance_type, self.max_run = instance_type, max_run
        self.model_dir, self.output_path = model_dir, output_path
        self.environment = dict(environment or {})
        self.base_job_name, self.python = base_job_name, python
        self.ignored = ignored              # role, image_uri, framework_version, py_version, disable_profiler, subnets, ...
        self.processes_per_host = processes_per_host or self._pph_from_distribution()
        self.latest_job_name: Optional[str] = None
        self.log_path: Optional[str] = None
        self.metrics: Dict[str, List[float]] = {}

    # ------------------------------------------------------------------ helpers
    def _pph_from_distribution(self) -> Optional[int]:
        mpi = self.distribution.get("mpi") or {}
        return int(mpi["processes_per_host"]) if mpi.get("processes_per_host") else None

    def launcher_kind(self) -> str:
        d = self.distribution
        if (d.get("mpi") or {}).get("enabled") or ((d.get("smdistributed") or {}).get("dataparallel") or {}).get("enabled"):
            return "mpi"
        return "torchrun"

    def _nproc(self) -> int:
        if self.processes_per_host:
            return int(self.processes_per_host)
        try:
            import torch
            n = torch.cuda.device_count()
        except Exception:
            n = 0
        return max(1, n)

    def _script(self) -> str:
        p = os.path.join(self.source_dir, self.entry_point) if self.source_dir else self.entry_point
        if not os.path.isfile(p):
            raise FileNotFoundError(f"entry point {p} not found")
        return os.path.abspath(p)

    def command(self) -> List[str]:
        return [self.python, self._script(), *hyperparameters_to_argv(self.hyperparameters)]

    def _env(self, rank: int, world: int, port: int, inputs: Dict[str, str], model_dir: str, out_dir: str) -> Dict[str, str]:
        env = dict(os.environ)
        env.setdefault("NCCL_DEBUG", "INFO")
        env.update({k: str(v) for k, v in self.environment.items()})
        for k in ("RANK", "LOCAL_RANK", "WORLD_SIZE", "OMPI_COMM_WORLD_RANK", "OMPI_COMM_WORLD_SIZE", "OMPI_COMM_WORLD_LOCAL_RANK"):
            env.pop(k, None)
        env.update({"MASTER_ADDR": "127.0.0.1", "MASTER_PORT": str(port)})
        if self.launcher_kind() == "mpi":
            env.update({"OMPI_COMM_WORLD_RANK": str(rank), "OMPI_COMM_WORLD_SIZE": str(world), "OMPI_COMM_WORLD_LOCAL_RANK": str(rank),
                        "OMPI_COMM_WORLD_LOCAL_SIZE": str(world)})
        else:
            env.update({"RANK": str(rank), "LOCAL_RANK": str(rank), "WORLD_SIZE": str(world), "LOCAL_WORLD_SIZE": str(world)})
        env.update({"SM_MODEL_DIR": model_dir, "SM_OUTPUT_DATA_DIR": out_dir, "SM_NUM_GPUS": str(world), "SM_HOSTS": json.dumps(["algo-1"]),
                    "SM_CURRENT_HOST": "algo-1", "SM_HPS": json.dumps(self.hyperparameters), "SM_CHANNELS": json.dumps(sorted(inputs)),
                    "SM_USER_ARGS": json.dumps(hyperparameters_to_argv(self.hyperparameters)), "SM_USER_ENTRY_POINT": self.entry_point,
                    "SM_TRAINING_ENV": json.dumps({"job_name": self.latest_job_name, "hosts": ["algo-1"], "num_gpus": world})})
        for name, path in inputs.items():
            env["SM_CHANNEL_" + name.upper()] = path
        return env

    # ------------------------------------------------------------------ the job
    def fit(self, inputs: Union[None, str, Dict[str, str]] = None, wait: bool = True, logs: bool = True, job_name: Optional[str] = None) -> Dict[str, List[float]]:
        if isinstance(inputs, str):
            inputs = {"training": inputs}
        inputs = {k: os.path.abspath(str(v)) for k, v in (inputs or {}).items()}
        for k, v in inputs.items():
            if not os.path.isdir(v):
                raise FileNotFoundError(f"channel '{k}': {v} is not a local directory (S3 URIs are not fetched here)")
        self.latest_job_name = job_name or f"{self.base_job_name}-{time.strftime('%Y-%m-%d-%H-%M-%S')}"
        root = os.path.abspath(self.output_path or os.path.join(os.getcwd(), "local_jobs"))
        job_dir = os.path.join(root, self.latest_job_name)
        model_dir = os.path.abspath(self.model_dir) if self.model_dir else os.path.join(job_dir, "model")
        out_dir = os.path.join(job_dir, "output")
        for d in (model_dir, out_dir):
            os.makedirs(d, exist_ok=True)
        self.log_path = os.path.join(job_dir, "job.log")
        world, port = self._nproc(), _free_port()
        cmd = self.command()
        cwd = os.path.abspath(self.source_dir) if self.source_dir else None
        lock = threading.Lock()
        procs: List[subprocess.Popen] = []
        with open(self.log_path, "w") as logf:
            header = f"[{self.launcher_kind()}] {world} process(es): {' '.join(shlex.quote(c) for c in cmd)}"
            logf.write(header + "\n")
            if logs:
                print(header, flush=True)

            def pump(rank, proc):
                for line in proc.stdout:
                    with lock:
                        logf.write(line)
                        if logs:
                            sys.stdout.write(f"[1,{rank}]<stdout>:{line}" if world > 1 else line)

            threads = []
            for r in range(world):
                p = subprocess.Popen(cmd, cwd=cwd, env=self._env(r, world, port, inputs, model_dir, out_dir), stdout=subprocess.PIPE,
                                     stderr=subprocess.STDOUT, text=True, bufsize=1, start_new_session=True)
                procs.append(p)
                t = threading.Thread(target=pump, args=(r, p), daemon=True)
                t.start()
                threads.append(t)
            failed = self._wait(procs)
            for t in threads:
                t.join(timeout=10)
        with open(self.log_path) as f:
            self.metrics = extract_metrics(f.read(), self.metric_definitions)
        if failed is not None:
            raise TrainingJobError(f"training job {self.latest_job_name} failed: {failed} (log: {self.log_path})")
        self.model_data = model_dir
        return self.metrics

    def _wait(self, procs) -> Optional[str]:
        deadline = time.time() + self.max_run if self.max_run else None
        failed = None
        while True:
            codes = [p.poll() for p in procs]
            bad = [(i, c) for i, c in enumerate(codes) if c not in (None, 0)]
            if bad:
                failed = f"rank {bad[0][0]} exited with code {bad[0][1]}"
                break
            if all(c == 0 for c in codes):
                return None
            if deadline and time.time() > deadline:
                failed = f"max_run of {self.max_run} s exceeded"
                break
            time.sleep(0.05)
        for p in procs:                      # fail-stop: take the remaining ranks down (their own process groups only)
            if p.poll() is None:
                try:
                    os.killpg(p.pid, signal.SIGTERM)
                except (ProcessLookupError, PermissionError):
                    pass
        t_end = time.time() + 10
        for p in procs:
            try:
                p.wait(timeout=max(0.1, t_end - time.time()))
            except subprocess.TimeoutExpired:
                try:
                    os.killpg(p.pid, signal.SIGKILL)
                except (ProcessLookupError, PermissionError):
                    pass
        return failed

    def logs(self) -> str:
        with open(self.log_path) as f:
            return f.read()


PyTorch = LocalEstimator          # the class name the notebooks use (sagemaker.pytorch.PyTorch)


def extract_metrics(log: str, metric_definitions: List[Dict]) -> Dict[str, List[float]]:
    """apply SageMaker metric regexes (first capture group = value) line by line"""
    out: Dict[str, List[float]] = {}
    compiled = [(m["Name"], re.compile(m["Regex"])) for m in metric_definitions]
    for line in log.splitlines():
        for name, rx in compiled:
            m = rx.search(line)
            if m:
                try:
                    out.setdefault(name, []).append(float(m.group(1)))
                except (ValueError, IndexError):
                    pass
    return out
