"""BASELINE.json configs[1]: fixed step budget on 1xB200 against the CPU reference loss curve.

Runs the same optimisation on (a) the B200 path (B200Trainer.training_step, device-generated windows) and (b) the
CPU oracle (oracle/lstm_oracle.py) fed the very same windows copied from the device, from the same initial
parameters, and reports per-step relative loss differences and final validation/test metrics.

    python tools/loss_curve_parity.py [--loss mse] [--batch 256] [--steps 20] [--big-steps 2] [--out gpurun_out/x.json]
"""
from __future__ import annotations

import argparse
import json
import sys
import time
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "oracle"))

import lstm_oracle as O  # noqa: E402  (checker)
import masters_thesis_b200 as B  # noqa: E402


def run(loss, batch, steps, engine, seed=1234, val_windows=64):
    dev = "cuda:0"
    torch.manual_seed(seed)
    cls = {"mse": B.FinancialLstmMse, "nll": B.FinancialLstmNll, "combined": B.FinancialLstmCombined}[loss]
    model = cls(3, 64, 2, 0.0, 1e-4)
    init = model.flat_parameters().clone()
    dm = B.FinancialLstmDataModule(lookback_window=60, target_window=30, stride=90, batch_size=batch, seed=seed, device=dev,
                                   streaming=True, steps_per_epoch=steps)
    dm.prepare_data(); dm.setup("fit")
    tr = B.B200Trainer(max_epochs=1, gradient_clip_val=5.0, engine=engine, verbose=False)
    tr._attach(model, dm)
    om = O.OracleLstm(3, 64, 2, 0.0)
    off = 0
    with torch.no_grad():
        for p in om.parameters():
            p.copy_(init[off:off + p.numel()].reshape(p.shape)); off += p.numel()
    opt = O.make_optimizer(om)
    rows = []
    t_gpu = t_cpu = 0.0
    for k in range(steps):
        b = dm.source.generate(first_window=dm.n_windows + k * batch, n_windows=batch)
        ref = [t.cpu() for t in b.to_reference()]
        torch.cuda.synchronize(); t0 = time.perf_counter()
        lg = float(tr.training_step(model, b, 1e-4)[2])
        torch.cuda.synchronize(); t_gpu += time.perf_counter() - t0
        t0 = time.perf_counter()
        lc, _ = O.train_step(om, opt, ref, loss, 5.0, dense_nll=(batch * 100 <= 3200))
        t_cpu += time.perf_counter() - t0
        rows.append({"step": k, "b200": lg, "cpu": lc["total"], "rel": abs(lg - lc["total"]) / abs(lc["total"])})
    vb = dm.source.generate(first_window=dm.train_end, n_windows=val_windows)
    vref = [t.cpu() for t in vb.to_reference()]
    v_gpu = [float(x) for x in model.engine.eval_windows(vb, "combined")]
    v_cpu = O.eval_windows(om, *vref, "combined")
    t_cpu_test = O.test_windows(om, *vref)
    dp = float((model.flat_parameters() - O.flat_params(om)).abs().max())
    return {"loss": loss, "batch_windows": batch, "steps": steps, "engine": engine, "curve": rows,
            "max_rel_loss_diff": max(r["rel"] for r in rows),
            "val": {"b200": {"mse": v_gpu[0], "nll": v_gpu[1], "total": v_gpu[2], "mae": v_gpu[3]},
                    "cpu": {**v_cpu, "mae": t_cpu_test["mae"]}},
            "max_abs_param_diff": dp, "seconds": {"b200": t_gpu, "cpu": t_cpu}}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--loss", default="mse,combined")
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--big-batch", type=int, default=4096)
    ap.add_argument("--big-steps", type=int, default=1)
    ap.add_argument("--engine", default="tc_precise")
    ap.add_argument("--out", default="gpurun_out/loss_curve.json")
    a = ap.parse_args()
    torch.set_float32_matmul_precision("highest")
    out = []
    for loss in a.loss.split(","):
        out.append(run(loss, a.batch, a.steps, a.engine))
        print(loss, "max rel loss diff", out[-1]["max_rel_loss_diff"], "param diff", out[-1]["max_abs_param_diff"], out[-1]["seconds"], flush=True)
    if a.big_steps > 0:
        out.append(run("mse", a.big_batch, a.big_steps, a.engine, val_windows=8))
        print("mse big", out[-1]["max_rel_loss_diff"], out[-1]["seconds"], flush=True)
    Path(a.out).parent.mkdir(parents=True, exist_ok=True)
    Path(a.out).write_text(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
