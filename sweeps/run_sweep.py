"""Runs a Hydra-style multirun grid as parallel single-GPU jobs (BASELINE.json configs[4]).

The reference's sweeps call `python3 train.py -m a=1,2 b=x,y` and rely on the joblib launcher with n_jobs: 1
(configs/config.yaml:17-20), i.e. one job after the other.  Here every job of the grid is an independent
`python train.py <overrides>` process pinned to one GPU with CUDA_VISIBLE_DEVICES = job index mod n_gpus, at most one
job per GPU at a time ("replicas only": no collective, no shared state).

    python sweeps/run_sweep.py [--gpus 8] [--dry-run] -- datamodule=synthetic model=small,medium,large loss=mse,nll,combined trainer=b200
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from masters_thesis_b200.config import expand_multirun  # noqa: E402


def plan(overrides, n_gpus):
    return [{"job": i, "gpu": i % n_gpus, "overrides": ov} for i, ov in enumerate(expand_multirun(overrides))]


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=int(os.environ.get("B200_SWEEP_GPUS", "8")))
    ap.add_argument("--dry-run", action="store_true")
    ap.add_argument("--log-dir", default="logs/sweeps")
    ap.add_argument("overrides", nargs="*")
    a = ap.parse_args(argv)
    jobs = plan([o for o in a.overrides if "=" in o], a.gpus)
    if a.dry_run:
        print(json.dumps(jobs, indent=1))
        return jobs
    log_dir = Path(a.log_dir)
    log_dir.mkdir(parents=True, exist_ok=True)
    running, results, queue = {}, [], list(jobs)
    t0 = time.time()
    while queue or running:
        free = [g for g in range(a.gpus) if g not in running]
        while queue and free:
            # keep the job -> GPU assignment deterministic: a job only starts on its own GPU
            nxt = next((j for j in queue if j["gpu"] in free), None)
            if nxt is None:
                break
            queue.remove(nxt)
            free.remove(nxt["gpu"])
            env = dict(os.environ, CUDA_VISIBLE_DEVICES=str(nxt["gpu"]))
            log = open(log_dir / f"job{nxt['job']:03d}.log", "w")
            p = subprocess.Popen([sys.executable, str(ROOT / "train.py"), *nxt["overrides"]], cwd=ROOT, env=env,
                                 stdout=log, stderr=subprocess.STDOUT)
            running[nxt["gpu"]] = (p, nxt, log, time.time())
        for g, (p, job, log, ts) in list(running.items()):
            rc = p.poll()
            if rc is not None:
                log.close()
                results.append(dict(job, returncode=rc, seconds=time.time() - ts))
                del running[g]
        time.sleep(0.2)
    summary = {"jobs": sorted(results, key=lambda r: r["job"]), "wall_seconds": time.time() - t0, "gpus": a.gpus}
    (log_dir / "summary.json").write_text(json.dumps(summary, indent=1))
    print(json.dumps(summary))
    return 0 if all(r["returncode"] == 0 for r in results) else 1


if __name__ == "__main__":
    sys.exit(main() or 0)
