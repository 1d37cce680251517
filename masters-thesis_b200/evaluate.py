"""Device-side mirror of the reference's evaluation script (test.py:14-89): for every test window the model's
(alpha, beta), the OLS (alpha, beta) of each stock on the market over the LOOK-BACK days, and their residuals
against the target returns and the true coefficients.  Many windows per launch; per-window semantics.

The reference draws scatter plots and histograms with matplotlib into TensorBoard (test.py:91-145, src/plots.py);
matplotlib is not installed here, so `write_results` logs the same quantities as TensorBoard histograms and scalars
and saves the tensors (`test_results.npz`) for plotting elsewhere.
"""
from __future__ import annotations

from pathlib import Path

import numpy as np
import torch


def lookback_ols(batch):
    """OLS of each stock's look-back returns on the market's (test.py:52 -> src/common.py:5-47): closed form of
    pinv(X^T X) X^T y for the design [1, r_m].  Returns (alpha, beta), each [N]."""
    rs, rm = batch.x[:, :, 0].double(), batch.x[:, :, 1].double()           # [T][N]
    ms, mm = rs.mean(0), rm.mean(0)
    beta = ((rm - mm) * (rs - ms)).sum(0) / ((rm - mm) ** 2).sum(0)
    return (ms - beta * mm).float(), beta.float()


@torch.no_grad()
def collect_test_results(model, dm, batch_windows=256, limit_windows=None):
    """The reference's `test_results` dictionary after transform_test_results (test.py:75-89): every entry [Nw][S]."""
    e = model.engine
    assert e is not None, "attach the model first"
    lo, hi = dm.val_end, dm.n_windows
    if limit_windows:
        hi = min(hi, lo + int(limit_windows))
    coeff = dm.source.coefficients()
    keys = [("recon_residuals", ("model", "ols")), ("alpha_residuals", ("model", "ols")), ("beta_residuals", ("model", "ols")),
            ("alpha", ("model", "ols", "true")), ("beta", ("model", "ols", "true"))]
    out = {k: {s: [] for s in sub} for k, sub in keys}
    done = 0
    while done < hi - lo:
        nb = min(batch_windows, hi - lo - done)
        b = dm.source.generate(first_window=lo + done, n_windows=nb)
        S, N = b.n_stocks, nb * b.n_stocks
        e.forward(b, training=False)
        e.heads(b, "mse", want_grad=False, windows_per_group=1)
        a_m, b_m = e.alpha[:N].clone(), e.beta[:N].clone()
        a_o, b_o = lookback_ols(b)
        if coeff is not None:                               # y channels 2, 3 = true per-stock values (src/data.py:213-214)
            a_t, b_t = coeff[0].repeat(nb), coeff[1].repeat(nb)
        else:                                               # ... or the per-window OLS of the target (src/data.py:205-207)
            a_t, b_t = b.ols_alpha[:N].clone(), b.ols_beta[:N].clone()
        rmt = b.y_market.repeat_interleave(S, dim=0)        # [N][n]
        rec_m = (b.y_stock - (a_m[:, None] + b_m[:, None] * rmt)).mean(-1)
        rec_o = (b.y_stock - (a_o[:, None] + b_o[:, None] * rmt)).mean(-1)
        for k, s, v in (("recon_residuals", "model", rec_m), ("recon_residuals", "ols", rec_o),
                        ("alpha_residuals", "model", a_t - a_m), ("alpha_residuals", "ols", a_t - a_o),
                        ("beta_residuals", "model", b_t - b_m), ("beta_residuals", "ols", b_t - b_o),
                        ("alpha", "model", a_m), ("alpha", "ols", a_o), ("alpha", "true", a_t),
                        ("beta", "model", b_m), ("beta", "ols", b_o), ("beta", "true", b_t)):
            out[k][s].append(v.reshape(nb, S).cpu())
        done += nb
    return {k: {s: torch.cat(v) for s, v in sub.items()} for k, sub in out.items()}


def write_results(results, log_dir, writer=None):
    """test_results.npz + TensorBoard histograms/scalars under the reference's tag families (hist/*, estimation/*)."""
    log_dir = Path(log_dir)
    log_dir.mkdir(parents=True, exist_ok=True)
    flat = {f"{k}/{s}": v.numpy() for k, sub in results.items() for s, v in sub.items()}
    np.savez_compressed(log_dir / "test_results.npz", **{k.replace("/", "__"): v for k, v in flat.items()})
    summary = {}
    for fam, tag in (("recon_residuals", "recon_residuals"), ("alpha_residuals", "alphas"), ("beta_residuals", "betas")):
        for s in ("model", "ols"):
            v = results[fam][s]
            summary[f"hist/{tag}/{s}_mean_abs"] = float(v.abs().mean())
            summary[f"hist/{tag}/{s}_std"] = float(v.std())
            if writer:
                writer.add_histogram(f"hist/{tag}/{s}", v.reshape(-1), 0)
    for kind in ("alpha", "beta"):
        for s in ("model", "ols"):
            summary[f"estimation/{kind}/{s}_rmse"] = float(((results[kind][s] - results[kind]["true"]) ** 2).mean().sqrt())
    if writer:
        for k, v in summary.items():
            writer.add_scalar(k, v, 0)
        writer.flush()
    return summary
