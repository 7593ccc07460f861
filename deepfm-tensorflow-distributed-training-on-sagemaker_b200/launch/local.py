"""Single-box launcher: what the two SageMaker notebooks do for the reference (NB-PS:56-143, NB-HVD:56-156):
turn a `hyperparameters` dict into `--key value` argv, provision `SM_*` environment variables, start one process per
worker, wire input channels in File or Pipe mode, wait for the job.

    est = LocalEstimator(hyperparameters={...}, distribution={"mpi": {"enabled": True, "processes_per_host": 8}},
                         input_mode="File")
    est.fit({"training": "/data/train", "evaluation": "/data/val"})

  distribution {"mpi": ...}               -> --strategy dp, N processes (one per GPU), NCCL/NVLink rendezvous env
  distribution {"parameter_server": ...}  -> --strategy ps, N worker processes sharing host-resident tables
  input_mode "Pipe"                        -> per-channel, per-epoch FIFOs <base>/<channel>_<epoch> fed by threads

CLI:  python -m deepfm_b200.launch --nproc_per_node 8 [--pipe] [--ps] -- --task_type train --field_size 39 ...
"""
from __future__ import annotations

import json
import os
import socket
import subprocess
import sys
import tempfile
import threading
import time
from typing import Dict, List, Optional


def _free_port() -> int:
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def hyperparameters_to_argv(hp: Dict[str, object]) -> List[str]:
    """SageMaker convention (NB-PS:89-94): every key becomes `--key value`; booleans become True/False strings."""
    argv: List[str] = []
    for k, v in hp.items():
        argv += ["--%s" % k, json.dumps(v) if isinstance(v, (list, dict)) else str(v)]
    return argv


class LocalEstimator:
    def __init__(self, hyperparameters: Dict[str, object], distribution: Optional[dict] = None, input_mode: str = "File",
                 instance_count: int = 1, entry_module: str = "deepfm_b200", python: str = sys.executable,
                 env: Optional[Dict[str, str]] = None, cwd: Optional[str] = None, max_restarts: int = 0,
                 grace_secs: float = 10.0):
        self.hp = dict(hyperparameters)
        self.dist = distribution or {}
        self.input_mode = input_mode
        self.instance_count = instance_count  # one box only; kept for API parity
        self.entry, self.python = entry_module, python
        self.extra_env = env or {}
        self.cwd = cwd
        mpi = self.dist.get("mpi", {})
        ps = self.dist.get("parameter_server", {})
        self.nproc = int(mpi.get("processes_per_host", 1)) if mpi.get("enabled") else \
            int(ps.get("workers_per_host", 2)) if ps.get("enabled") else 1
        if ps.get("enabled"):
            self.hp.setdefault("strategy", "ps")
        self.returncodes: List[int] = []
        self.max_restarts, self.grace_secs, self.restarts = int(max_restarts), float(grace_secs), 0

    def fit(self, inputs: Dict[str, str], wait: bool = True, timeout: Optional[float] = None) -> List[int]:
        pipe = self.input_mode.lower() == "pipe"
        hp = dict(self.hp)
        hp.setdefault("worker_per_host", self.nproc)
        base = None
        feeders: List[threading.Thread] = []
        if pipe:
            hp["pipe_mode"] = 1
            base = tempfile.mkdtemp(prefix="dfm_pipe_")
            hp["data_base_dir"] = base
            epochs = int(hp.get("num_epochs", 10))
            for ch, src in inputs.items():
                n_ep = 1 if ch.startswith("eval") else epochs
                for e in range(n_ep):
                    fifo = os.path.join(base, "%s_%d" % (ch, e))
                    os.mkfifo(fifo)
                    t = threading.Thread(target=_feed_fifo, args=(fifo, src), daemon=True)
                    t.start()
                    feeders.append(t)
        else:
            if "training" in inputs:
                hp.setdefault("training_data_dir", inputs["training"])
            if "evaluation" in inputs:
                hp.setdefault("val_data_dir", inputs["evaluation"])
        argv = hyperparameters_to_argv(hp)
        attempt = 0
        while True:
            procs = self._spawn(argv, inputs)
            if not wait:
                self._procs = procs
                return []
            try:
                self.returncodes = self._supervise(procs, timeout)
            except BaseException:  # Ctrl-C, timeout, ...: never leave workers (and their GPUs) behind
                for p in procs:
                    if p.poll() is None:
                        p.kill()
                raise
            if not any(self.returncodes):
                return self.returncodes
            # A failed attempt is restarted like an interrupted (spot) SageMaker job: same model_dir, so the chief
            # restores the latest checkpoint (Estimator semantics). Pipe mode cannot be replayed (the
            # FIFOs of the failed attempt are consumed), so only File mode restarts.
            if attempt >= self.max_restarts or pipe:
                raise RuntimeError("training job failed: return codes %s" % self.returncodes)
            attempt += 1
            self.restarts = attempt
            print("INFO:deepfm_b200.launch: attempt %d failed with return codes %s; restarting (%d/%d)"
                  % (attempt, self.returncodes, attempt, self.max_restarts), flush=True)

    def _spawn(self, argv: List[str], inputs: Dict[str, str]) -> List[subprocess.Popen]:
        port = _free_port()
        procs = []
        for r in range(self.nproc):
            env = dict(os.environ)
            env.update(self.extra_env)
            env.update({"RANK": str(r), "LOCAL_RANK": str(r), "WORLD_SIZE": str(self.nproc), "MASTER_ADDR": "127.0.0.1",
                        "MASTER_PORT": str(port), "SM_HOSTS": json.dumps(["algo-1"]), "SM_CURRENT_HOST": "algo-1",
                        "SM_CHANNELS": json.dumps(sorted(inputs)), "SM_NUM_GPUS": str(self.nproc)})
            for ch, src in inputs.items():
                env["SM_CHANNEL_%s" % ch.upper().replace("-", "_")] = src
            procs.append(subprocess.Popen([self.python, "-m", self.entry] + argv, env=env, cwd=self.cwd))
        return procs

    def _supervise(self, procs: List[subprocess.Popen], timeout: Optional[float]) -> List[int]:
        """Failure detection (the reference relies on Horovod's 'shut down because a rank exited' error, PDF p22): poll
        every worker; when one exits non-zero the survivors - which would otherwise sit in a collective until the
        device-side watchdog fires - get `grace_secs` to notice by themselves and are then terminated."""
        t0 = time.time()
        failed_at: Optional[float] = None
        while True:
            codes = [p.poll() for p in procs]
            if all(c is not None for c in codes):
                return [int(c) for c in codes]
            now = time.time()
            if failed_at is None and any(c not in (None, 0) for c in codes):
                failed_at = now
            if failed_at is not None and now - failed_at > self.grace_secs:
                for p in procs:
                    if p.poll() is None:
                        p.terminate()
                deadline = time.time() + 5.0
                for p in procs:
                    try:
                        p.wait(timeout=max(0.1, deadline - time.time()))
                    except subprocess.TimeoutExpired:
                        p.kill()
                        p.wait()
                return [int(p.returncode) for p in procs]
            if timeout is not None and now - t0 > timeout:
                for p in procs:
                    if p.poll() is None:
                        p.kill()
                raise subprocess.TimeoutExpired(cmd="%s -m %s" % (self.python, self.entry), timeout=timeout)
            time.sleep(0.05)


def _feed_fifo(fifo: str, src: str) -> None:
    """Stream every *.tfrecords file under `src` (or the file `src`) into the FIFO, like SageMaker's pipe agent."""
    import glob
    files = [src] if os.path.isfile(src) else sorted(glob.glob(os.path.join(src, "**", "*.tfrecords"), recursive=True))
    try:
        with open(fifo, "wb") as out:
            for f in files:
                with open(f, "rb") as fin:
                    while True:
                        chunk = fin.read(1 << 20)
                        if not chunk:
                            break
                        out.write(chunk)
    except BrokenPipeError:
        pass


def main(argv=None) -> int:
    import argparse
    argv = list(sys.argv[1:] if argv is None else argv)
    rest: List[str] = []
    if "--" in argv:
        i = argv.index("--")
        argv, rest = argv[:i], argv[i + 1:]
    ap = argparse.ArgumentParser(description="single-box launcher (stands in for the SageMaker notebooks)")
    ap.add_argument("--nproc_per_node", type=int, default=1)
    ap.add_argument("--ps", action="store_true", help="parameter-server distribution instead of mpi/dp")
    ap.add_argument("--pipe", action="store_true", help="Pipe input mode")
    ap.add_argument("--training", default="")
    ap.add_argument("--evaluation", default="")
    ap.add_argument("--max_restarts", type=int, default=0,
                    help="restart a failed File-mode job from its latest checkpoint up to this many times")
    a = ap.parse_args(argv)
    hp: Dict[str, object] = {}
    k = None
    for tok in rest:  # pass-through flags -> hyperparameters
        if tok.startswith("--"):
            k = tok[2:]
            hp[k] = "True"
        elif k is not None:
            hp[k] = tok
            k = None
    dist = {"parameter_server": {"enabled": True, "workers_per_host": a.nproc_per_node}} if a.ps else \
        {"mpi": {"enabled": True, "processes_per_host": a.nproc_per_node}}
    inputs = {}
    if a.training:
        inputs["training"] = a.training
    if a.evaluation:
        inputs["evaluation"] = a.evaluation
    LocalEstimator(hp, dist, "Pipe" if a.pipe else "File", max_restarts=a.max_restarts).fit(inputs)
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
