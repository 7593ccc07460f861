"""Manual cluster-spec builder: the role taxonomy of the reference's `set_dist_env` (PS:294-339), which is dead code
there (never called, PS:388, and it reads an undefined flag, PS:297) but documents the intended roles:

    dist_mode 1 (local test)   : 1 chief + 1 ps (+ optional evaluator)
    dist_mode 2 (multi-worker) : worker_hosts[0] -> chief, worker_hosts[1] -> evaluator, the rest -> workers; ps_hosts -> ps

Returned in TF_CONFIG form (the dict SageMaker would have put in the environment, README-EN:16) and mapped onto this
framework's processes: chief/worker -> training ranks (chief = rank 0), ps -> the shared host store (no process),
evaluator -> a process running `--task_type eval` against the same model_dir.
"""
from __future__ import annotations

import json
from typing import Dict, List, Optional


def _split(s) -> List[str]:
    if isinstance(s, (list, tuple)):
        return list(s)
    return [x for x in str(s).split(",") if x]


def build_tf_config(dist_mode: int, ps_hosts, worker_hosts, job_name: str, task_index: int,
                    chief_hosts=None) -> Optional[Dict]:
    ps, workers, chiefs = _split(ps_hosts), _split(worker_hosts), _split(chief_hosts or "")
    if dist_mode == 1:
        chief = chiefs or workers[:1]
        cluster = {"chief": chief, "ps": ps}
        task = {"type": job_name, "index": task_index}
    elif dist_mode == 2:
        if not workers:
            raise ValueError("dist_mode 2 needs worker_hosts")
        cluster = {"chief": workers[:1], "ps": ps}
        if len(workers) > 2:
            cluster["worker"] = workers[2:]
        if job_name == "worker" and task_index == 0:
            task = {"type": "chief", "index": 0}
        elif job_name == "worker" and task_index == 1:
            task = {"type": "evaluator", "index": 0}
        elif job_name == "worker":
            task = {"type": "worker", "index": task_index - 2}
        else:
            task = {"type": job_name, "index": task_index}
    else:
        return None
    return {"cluster": cluster, "task": task}


def role_to_rank(tf_config: Dict) -> Dict:
    """Map a TF_CONFIG role onto this framework: training rank (chief=0, workers 1..), or a non-training role."""
    t = tf_config["task"]
    n_workers = len(tf_config["cluster"].get("worker", []))
    if t["type"] == "chief":
        return {"role": "train", "rank": 0, "world_size": 1 + n_workers}
    if t["type"] == "worker":
        return {"role": "train", "rank": 1 + t["index"], "world_size": 1 + n_workers}
    if t["type"] == "evaluator":
        return {"role": "eval"}
    return {"role": "store"}  # ps: parameters live in the shared host store, no process needed


def tf_config_env(tf_config: Optional[Dict]) -> Dict[str, str]:
    return {} if tf_config is None else {"TF_CONFIG": json.dumps(tf_config)}
