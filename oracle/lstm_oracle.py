"""CPU oracle for the training hot path -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this module; the product path (masters-thesis_b200/) never does
and fails loudly when its CUDA library is missing.

This is a restatement (fresh code, same arithmetic) of the reference's per-step
training loop on torch CPU fp32 -- torch is the library the reference itself
computes with, so the restatement uses the same aten kernels (oneDNN LSTM,
linalg.pinv, logdet) the reference would run on the host:

    data prep      /root/reference/src/common.py:81-149, src/data.py:126-130
    model          /root/reference/src/model.py:77-109
    objectives     /root/reference/src/model.py:12-69, 192-202, 234-249, 286-306
                   /root/reference/src/common.py:50-78
    optimiser      /root/reference/src/model.py:149-173, train.py:171
    eval steps     /root/reference/src/model.py:119-147, 211-215, 258-262, 321-331

Third-party behaviour that is restated: Lightning's step
order (forward -> zero_grad -> backward -> clip_grad_norm_ -> Adam.step),
torchmetrics' "value of this batch" forward, DataLoader collation.

PINNING: the reference has no tests or golden vectors.  The
restatement is pinned against the reference's own modules, imported unmodified
behind oracle/ref_stubs.py in the build container; oracle/make_golden.py writes
their outputs to tests/golden/*.npz and tests/test_oracle.py checks this file
against those vectors.

The second half of the file is the written specification of the device batch
generator (Philox4x32-10 -> Box-Muller -> Student-t(5) single-factor DGP,
restating the distributions of /root/reference/src/data.py:36-59) in numpy.
Integer Philox output is compared bit-exactly with the CUDA kernel; the float
transform within a few ulp.
"""
from __future__ import annotations

import math
from dataclasses import dataclass

import numpy as np
import torch

LOG_2PI = math.log(2.0 * math.pi)

# --------------------------------------------------------------------------- data prep


def split_windows(r_stocks, r_market, lookback=60, target=30, stride=None, prediction=True):
    """(S, D) stock and (D,) market series -> X (Nw, S, lookback, 2), y (Nw, S, target, 2).

    Follows lookback_target_split (src/common.py:81-112): channel 0 is the stock return, channel 1 the
    market return broadcast over stocks.  prediction=True: windows of lookback+target days start every
    `stride` days and y is the part after the look-back; prediction=False (reconstruction,
    src/common.py:107-110): windows are `lookback` days long and y is their last `target` days.
    """
    if stride is None:
        stride = lookback + target
    S, D = r_stocks.shape
    total = lookback + target if prediction else lookback
    n_win = (D - total) // stride + 1
    starts = torch.arange(n_win) * stride
    idx = starts[:, None] + torch.arange(total)[None, :]          # (Nw, total)
    stock = r_stocks[:, idx].permute(1, 0, 2)                       # (Nw, S, total)
    market = r_market[idx][:, None, :].expand(-1, S, -1)            # (Nw, S, total)
    win = torch.stack([stock, market], dim=-1)                      # (Nw, S, total, 2)
    if prediction:
        return win[:, :, :lookback, :], win[:, :, lookback:, :]
    return win, win[:, :, lookback - target:, :]


def quadratic_features(X, interaction_only=True):
    """[r_s, r_m, r_s*r_m] (+ r_s^2, r_m^2) -- src/common.py:115-130."""
    rs, rm = X[..., 0], X[..., 1]
    feats = [rs, rm, rs * rm]
    if not interaction_only:
        feats += [rs * rs, rm * rm]
    return torch.stack(feats, dim=-1)


def ols_targets(y):
    """Per-window OLS of every stock on the market over the TARGET days.

    src/common.py:132-149 with _batched_ols (src/common.py:33-47): the normal equations
    are solved through pinv(X^T X) in fp32, residual and market variances are the
    unbiased ones (torch.var default).  Returns alphas, betas (Nw, S), factor (Nw, 2)
    = [mean, var] of the market, inv_psi (Nw, S).
    """
    rs = y[:, :, :, 0]                                  # (Nw, S, n)
    rm = y[:, 0, :, 1]                                  # (Nw, n)
    design = torch.stack([torch.ones_like(rm), rm], dim=-1)            # (Nw, n, 2)
    gram = design.mT @ design                                            # (Nw, 2, 2)
    coef = torch.linalg.pinv(gram) @ (design.mT @ rs.mT)                 # (Nw, 2, S)
    alphas, betas = coef[:, 0, :], coef[:, 1, :]
    resid = rs - (alphas[..., None] + betas[..., None] * rm[:, None, :])
    factor = torch.stack([rm.mean(-1), rm.var(-1)], dim=-1)
    inv_psi = 1.0 / resid.var(-1)
    return alphas, betas, factor, inv_psi


def build_dataset(r_stocks, r_market, lookback=60, target=30, stride=90, interaction_only=True,
                  alphas=None, betas=None, prediction=True):
    """The four tensors of the reference TensorDataset (src/data.py:193-216).

    X (Nw,S,60,F), y (Nw,S,30,4) with the true alpha/beta appended as channels 2,3
    (src/data.py:126-130, 213-214), factor (Nw,2), inv_psi (Nw,S).
    """
    X, y = split_windows(r_stocks, r_market, lookback, target, stride, prediction)
    X = quadratic_features(X, interaction_only)
    a_hat, b_hat, factor, inv_psi = ols_targets(y)
    if alphas is None or betas is None:
        # real data: per-window OLS estimates (Nw, S)
        a_ch, b_ch = a_hat[:, :, None, None], b_hat[:, :, None, None]
    else:
        a_ch, b_ch = alphas.view(1, -1, 1, 1), betas.view(1, -1, 1, 1)
    Nw, S, n, _ = y.shape
    y = torch.cat([y, a_ch.expand(Nw, S, n, 1), b_ch.expand(Nw, S, n, 1)], dim=-1)
    return X.contiguous(), y.contiguous(), factor, inv_psi


def dataset_split(n):
    """70/20/10 contiguous split (src/data.py:221-234): (train_end, val_end)."""
    return int(0.7 * n), int(0.9 * n)


# --------------------------------------------------------------------------- model


class OracleLstm(torch.nn.Module):
    """nn.LSTM encoder + alpha/beta heads, src/model.py:77-109.

    Attribute names (and therefore state_dict keys and the order in which the
    default initialisers consume the torch RNG) match the reference module.
    """

    def __init__(self, input_size=3, hidden_size=64, num_layers=2, dropout=0.2):
        super().__init__()
        self.lstm = torch.nn.LSTM(input_size=input_size, hidden_size=hidden_size,
                                  num_layers=num_layers,
                                  dropout=dropout if num_layers > 1 else 0,
                                  batch_first=True)
        self.alpha_head = torch.nn.Linear(hidden_size, 1)
        self.beta_head = torch.nn.Linear(hidden_size, 1)

    def forward(self, x):
        out, _ = self.lstm(x)
        last = out[:, -1, :]
        return self.alpha_head(last), self.beta_head(last)


class ExplicitMaskLstm(torch.nn.Module):
    """Same network as OracleLstm, built from single-layer nn.LSTMs so that the
    inter-layer dropout masks can be injected (shape (L-1, N, T, H), already
    scaled by 1/(1-p); indexed masks[l] = mask applied to the OUTPUT of layer l).  Used for the shared-Philox-mask parity gate and to read
    every layer's hidden sequence."""

    def __init__(self, src: OracleLstm):
        super().__init__()
        L, H = src.lstm.num_layers, src.lstm.hidden_size
        self.layers = torch.nn.ModuleList()
        for l in range(L):
            cell = torch.nn.LSTM(src.lstm.input_size if l == 0 else H, H, 1, batch_first=True)
            for name in ("weight_ih", "weight_hh", "bias_ih", "bias_hh"):
                # share the Parameter objects, so gradients land on the source module
                setattr(cell, f"{name}_l0", getattr(src.lstm, f"{name}_l{l}"))
            self.layers.append(cell)
        self.alpha_head, self.beta_head = src.alpha_head, src.beta_head

    def forward(self, x, masks=None, return_all=False):
        seqs = []
        for l, cell in enumerate(self.layers):
            x, _ = cell(x)
            seqs.append(x)
            if masks is not None and l < len(self.layers) - 1:
                x = x * masks[l]
        last = x[:, -1, :]
        out = (self.alpha_head(last), self.beta_head(last))
        return out + (seqs,) if return_all else out


# --------------------------------------------------------------------------- objectives


def woodbury_inverse_cov(beta, inv_psi_diag, f_var):
    """Sigma^-1 = Psi^-1 - (Psi^-1 b)(b^T Psi^-1)/(1/s2 + b^T Psi^-1 b), dense (K,K).
    src/common.py:50-78."""
    pb = inv_psi_diag @ beta                       # (K,1)
    bp = beta.T @ inv_psi_diag                     # (1,K)
    denom = 1.0 / f_var + bp @ beta                # (1,1)
    return inv_psi_diag - (pb @ bp) / denom


def gaussian_nll_dense(mean, inv_cov, target):
    """0.5*[n*(K*ln2pi - logdet(Sigma^-1)) + tr(D^T Sigma^-1 D)] -- src/model.py:54-69."""
    K, n = target.shape
    d = target - mean
    quad = torch.trace(d.T @ inv_cov @ d)
    return 0.5 * (n * (K * LOG_2PI - torch.logdet(inv_cov)) + quad)


def gaussian_nll_closed_form(alpha, beta, inv_psi, f_mean, f_var, target):
    """The same value in O(K*n) using the diagonal-plus-rank-1 structure (SURVEY.md
    section 8 a10).  alpha, beta, inv_psi: (K,), target (K, n).  Any dtype; used in
    fp64 as the large-batch oracle, validated against the dense path for K<=3200."""
    K, n = target.shape
    d = target - (alpha + beta * f_mean)[:, None]
    s = torch.sum(inv_psi * beta * beta)
    u = torch.sum((inv_psi * beta)[:, None] * d, dim=0)            # (n,)
    logdet = torch.sum(torch.log(inv_psi)) - torch.log1p(f_var * s)
    quad = torch.sum(inv_psi[:, None] * d * d) - torch.sum(u * u) / (1.0 / f_var + s)
    return 0.5 * (n * (K * LOG_2PI - logdet) + quad)


def unpack_batch(batch):
    """flatten(0,1) of the collated batch, as every _shared_step does
    (src/model.py:193-194, 235-238, 287-290)."""
    X, y, factor, inv_psi = batch
    ctx = X.flatten(0, 1)
    tgt = y.flatten(0, 1)
    return ctx, tgt[:, :, 0], tgt[:, :, 1], factor.flatten(0, 1), inv_psi.flatten(0, 1)


def batch_losses(model, batch, kind, mse_weight=100.0, dense_nll=True, forward=None):
    """Losses of one collated batch: dict with 'total' and whichever of 'mse', 'nll'
    the objective computes.  kind in {'mse','nll','combined'}.

    NLL on a batch of B>1 windows is the reference's literal behaviour: one joint
    Gaussian over K=B*S rows using the FIRST window's (f_mean, f_var)
    (src/model.py:237-247)."""
    ctx, r_tgt, m_tgt, factor, inv_psi = unpack_batch(batch)
    alpha, beta = (forward or model)(ctx)
    out = {}
    if kind in ("mse", "combined"):
        pred = alpha + beta * m_tgt
        out["mse"] = torch.mean((pred - r_tgt) ** 2)        # MeanSquaredError batch value
    if kind in ("nll", "combined"):
        f_mean, f_var = factor[0], factor[1]
        if dense_nll:
            inv_cov = woodbury_inverse_cov(beta, torch.diag(inv_psi), f_var)
            out["nll"] = gaussian_nll_dense(alpha + beta * f_mean, inv_cov, r_tgt)
        else:
            out["nll"] = gaussian_nll_closed_form(alpha[:, 0], beta[:, 0], inv_psi,
                                                  f_mean, f_var, r_tgt)
    if kind == "mse":
        out["total"] = out["mse"]
    elif kind == "nll":
        out["total"] = out["nll"]
    else:
        out["total"] = out["nll"] + mse_weight * out["mse"]   # src/model.py:317
    return out


def make_optimizer(model, lr=1e-4, weight_decay=1e-5):
    """Adam with L2-coupled weight decay, src/model.py:150-154."""
    return torch.optim.Adam(model.parameters(), lr=lr, weight_decay=weight_decay)


def clip_global_norm(params, max_norm):
    """Lightning's gradient_clip_val (train.py:171) -> torch clip_grad_norm_: scale all
    grads by min(1, c/(||g||_2 + 1e-6)).  Returns the pre-clip norm."""
    return torch.nn.utils.clip_grad_norm_(params, max_norm)


def train_step(model, opt, batch, kind, clip, mse_weight=100.0, dense_nll=True, forward=None):
    """One optimisation step in Lightning's order.  Returns (losses, pre-clip grad norm)."""
    model.train()
    losses = batch_losses(model, batch, kind, mse_weight, dense_nll, forward)
    opt.zero_grad()
    losses["total"].backward()
    gnorm = clip_global_norm(list(model.parameters()), clip) if clip is not None else None
    opt.step()
    return {k: float(v.detach()) for k, v in losses.items()}, (None if gnorm is None else float(gnorm))


def flat_params(model):
    return torch.cat([p.detach().reshape(-1) for p in model.parameters()])


def flat_grads(model):
    return torch.cat([p.grad.detach().reshape(-1) for p in model.parameters()])


def ddp_mean_grads(model, rank_batches, kind, mse_weight=100.0, dense_nll=True):
    """What DDP would hand the optimiser: mean over ranks of the local-batch gradients
    (SURVEY.md section 8e).  Leaves the mean in p.grad; returns the per-rank losses."""
    acc, losses = None, []
    for b in rank_batches:
        model.zero_grad()
        ls = batch_losses(model, b, kind, mse_weight, dense_nll)
        ls["total"].backward()
        g = [p.grad.detach().clone() for p in model.parameters()]
        acc = g if acc is None else [a + x for a, x in zip(acc, g)]
        losses.append({k: float(v.detach()) for k, v in ls.items()})
    for p, a in zip(model.parameters(), acc):
        p.grad = a / len(rank_batches)
    return losses


@torch.no_grad()
def eval_windows(model, X, y, factor, inv_psi, kind, mse_weight=100.0):
    """validation loop: batch = 1 window, eval mode; returns the epoch means that Lightning
    logs as loss/{mse,nll,total}/val (src/model.py:211-215, 258-262, 321-331)."""
    model.eval()
    sums = {}
    n = X.shape[0]
    for w in range(n):
        b = (X[w:w + 1], y[w:w + 1], factor[w:w + 1], inv_psi[w:w + 1])
        ls = batch_losses(model, b, kind, mse_weight)
        for k, v in ls.items():
            sums[k] = sums.get(k, 0.0) + float(v)
    return {k: v / n for k, v in sums.items()}


@torch.no_grad()
def test_windows(model, X, y, factor, inv_psi):
    """test loop (src/model.py:119-147): MAE over all elements, NLL mean over windows."""
    model.eval()
    abs_sum, count, nll_sum = 0.0, 0, 0.0
    n = X.shape[0]
    for w in range(n):
        b = (X[w:w + 1], y[w:w + 1], factor[w:w + 1], inv_psi[w:w + 1])
        ctx, r_tgt, m_tgt, fac, ip = unpack_batch(b)
        alpha, beta = model(ctx)
        abs_sum += float(torch.sum(torch.abs(alpha + beta * m_tgt - r_tgt)))
        count += r_tgt.numel()
        inv_cov = woodbury_inverse_cov(beta, torch.diag(ip), fac[1])
        nll_sum += float(gaussian_nll_dense(alpha + beta * fac[0], inv_cov, r_tgt))
    return {"mae": abs_sum / count, "nll": nll_sum / n}


def lookback_ols(context):
    """ols(context[0, :, 1], context[:, :, 0]) of test.py:52 (src/common.py:5-47): per-stock OLS on the market over the
    look-back days through the same pinv(X^T X) X^T y route."""
    X = context[0, :, 1]
    y = context[:, :, 0]
    D = torch.stack([torch.ones_like(X), X], dim=-1)[None]                  # (1, T, 2)
    sol = torch.matmul(torch.linalg.pinv(torch.matmul(D.mT, D)), torch.matmul(D.mT, y[None].mT))
    return sol[0, 0, :], sol[0, 1, :]


@torch.no_grad()
def script_test_windows(model, X, y):
    """The evaluation script's per-window quantities (test.py:41-89, after transform_test_results): a dictionary of
    (Nw, S) tensors -- reconstruction residuals (mean over target days), alpha/beta of model, look-back OLS and
    truth (y channels 2, 3), and their residuals."""
    model.eval()
    out = {}

    def put(k, s, v):
        out.setdefault(k, {}).setdefault(s, []).append(v)
    for w in range(X.shape[0]):
        ctx, tgt = X[w], y[w]
        r_t, m_t, a_t, b_t = tgt[:, :, 0], tgt[:, :, 1], tgt[:, 0, 2], tgt[:, 0, 3]
        a_m, b_m = model(ctx)
        a_o, b_o = lookback_ols(ctx)
        put("recon_residuals", "model", (r_t - (a_m + b_m * m_t)).mean(-1))
        put("recon_residuals", "ols", (r_t - (a_o[:, None] + b_o[:, None] * m_t)).mean(-1))
        a_m, b_m = a_m.squeeze(), b_m.squeeze()
        for k, t_, m_, o_ in (("alpha", a_t, a_m, a_o), ("beta", b_t, b_m, b_o)):
            put(k + "_residuals", "model", t_ - m_)
            put(k + "_residuals", "ols", t_ - o_)
            put(k, "model", m_)
            put(k, "ols", o_)
            put(k, "true", t_)
    return {k: {s: torch.stack(v) for s, v in sub.items()} for k, sub in out.items()}


class PlateauSchedule:
    """ReduceLROnPlateau(mode='min', factor=0.5, patience=2) with torch defaults
    (threshold 1e-4 relative, cooldown 0, min_lr 0, eps 1e-8) -- src/model.py:156-161."""

    def __init__(self, lr, factor=0.5, patience=2, threshold=1e-4, eps=1e-8):
        self.lr, self.factor, self.patience, self.threshold, self.eps = lr, factor, patience, threshold, eps
        self.best, self.bad = math.inf, 0

    def step(self, metric):
        if metric < self.best * (1.0 - self.threshold):
            self.best, self.bad = metric, 0
        else:
            self.bad += 1
        if self.bad > self.patience:
            new_lr = self.lr * self.factor
            if self.lr - new_lr > self.eps:
                self.lr = new_lr
            self.bad = 0
        return self.lr


# --------------------------------------------------------------------------- generator spec

PHILOX_M0 = np.uint64(0xD2511F53)
PHILOX_M1 = np.uint64(0xCD9E8D57)
PHILOX_W0 = np.uint32(0x9E3779B9)
PHILOX_W1 = np.uint32(0xBB67AE85)

ID_MARKET = 0xFFFFFFFF      # counter word 2 for the market series
ID_COEFF = 0xFFFFFFFE       # counter word 2 for the per-stock alpha/beta draw

MKT_LOC, MKT_SCALE = 0.0678, 0.5099       # Student-t(5), src/data.py:36
IDIO_SCALE = 0.3140                        # Student-t(5), loc 0, src/data.py:37
ALPHA_LOC, ALPHA_SCALE = 0.0098, 0.1271    # Normal, src/data.py:38
BETA_LOC, BETA_SCALE = 0.9444, 0.3521      # Normal, src/data.py:39


def philox4x32_10(c0, c1, c2, c3, k0, k1):
    """Philox4x32 with 10 rounds on arrays of uint32 counters (Salmon et al., SC'11).
    Returns four uint32 arrays."""
    c0, c1, c2, c3 = (np.asarray(c, dtype=np.uint32).copy() for c in np.broadcast_arrays(c0, c1, c2, c3))
    k0 = np.uint32(k0)
    k1 = np.uint32(k1)
    for _ in range(10):
        p0 = PHILOX_M0 * c0.astype(np.uint64)
        p1 = PHILOX_M1 * c2.astype(np.uint64)
        hi0, lo0 = (p0 >> np.uint64(32)).astype(np.uint32), p0.astype(np.uint32)
        hi1, lo1 = (p1 >> np.uint64(32)).astype(np.uint32), p1.astype(np.uint32)
        c0, c1, c2, c3 = hi1 ^ c1 ^ k0, lo1, hi0 ^ c3 ^ k1, lo0
        with np.errstate(over="ignore"):
            k0 = np.uint32(k0 + PHILOX_W0)
            k1 = np.uint32(k1 + PHILOX_W1)
    return c0, c1, c2, c3


def _uniform(x):
    """uint32 -> float32 in (0,1): ((x >> 8) + 0.5) * 2^-24, exact in fp32."""
    return ((x >> np.uint32(8)).astype(np.float32) + np.float32(0.5)) * np.float32(2.0 ** -24)


def _box_muller(xa, xb):
    ua, ub = _uniform(xa), _uniform(xb)
    r = np.sqrt(np.float32(-2.0) * np.log(ua)).astype(np.float32)
    ang = (np.float32(2.0 * math.pi) * ub).astype(np.float32)
    return (r * np.cos(ang)).astype(np.float32), (r * np.sin(ang)).astype(np.float32)


def student_t5(day, ident, seed):
    """One Student-t(5) draw per (day, ident): Z0 / sqrt((Z1^2+..+Z5^2)/5) from two Philox
    calls (counter word 3 = 0, 1) and three Box-Muller pairs (normals 6,7 unused)."""
    day = np.asarray(day, dtype=np.uint64)
    lo, hi = (day & np.uint64(0xFFFFFFFF)).astype(np.uint32), (day >> np.uint64(32)).astype(np.uint32)
    k0, k1 = seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF
    a = philox4x32_10(lo, hi, np.uint32(ident) if np.isscalar(ident) else ident, np.uint32(0), k0, k1)
    b = philox4x32_10(lo, hi, np.uint32(ident) if np.isscalar(ident) else ident, np.uint32(1), k0, k1)
    z0, z1 = _box_muller(a[0], a[1])
    z2, z3 = _box_muller(a[2], a[3])
    z4, z5 = _box_muller(b[0], b[1])
    chi = (z1 * z1 + z2 * z2 + z3 * z3 + z4 * z4 + z5 * z5).astype(np.float32)
    return (z0 / np.sqrt(chi * np.float32(0.2))).astype(np.float32)


def stock_coefficients(n_stocks, seed):
    """alpha_i, beta_i ~ Normal, drawn once per stock per run (src/data.py:53-54)."""
    i = np.arange(n_stocks, dtype=np.uint32)
    k0, k1 = seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF
    x = philox4x32_10(i, np.uint32(0), np.uint32(ID_COEFF), np.uint32(0), k0, k1)
    z0, z1 = _box_muller(x[0], x[1])
    alpha = (np.float32(ALPHA_LOC) + np.float32(ALPHA_SCALE) * z0).astype(np.float32)
    beta = (np.float32(BETA_LOC) + np.float32(BETA_SCALE) * z1).astype(np.float32)
    return alpha, beta


def generate_days(day0, n_days, n_stocks, seed):
    """r_stocks (S, n_days), r_market (n_days,), alphas (S,), betas (S,) for absolute days
    [day0, day0+n_days): r = alpha + beta*r_m + eps (src/data.py:49-59)."""
    days = np.arange(day0, day0 + n_days, dtype=np.uint64)
    r_m = (np.float32(MKT_LOC) + np.float32(MKT_SCALE) * student_t5(days, ID_MARKET, seed)).astype(np.float32)
    ids = np.arange(n_stocks, dtype=np.uint32)[:, None]
    eps = (np.float32(IDIO_SCALE) * student_t5(days[None, :], ids, seed)).astype(np.float32)
    alpha, beta = stock_coefficients(n_stocks, seed)
    r_s = ((alpha[:, None] + beta[:, None] * r_m[None, :]).astype(np.float32) + eps).astype(np.float32)
    return r_s, r_m, alpha, beta


def generate_windows(window_idx, n_stocks=100, lookback=60, target=30, stride=90, seed=1234,
                     interaction_only=True):
    """Reference-format batch for the given global window indices, built from the
    generator spec + the data-prep restatement above."""
    Xs, ys, fs, ps = [], [], [], []
    for w in window_idx:
        r_s, r_m, a, b = generate_days(int(w) * stride, lookback + target, n_stocks, seed)
        X, y, f, p = build_dataset(torch.from_numpy(r_s), torch.from_numpy(r_m), lookback, target,
                                   lookback + target, interaction_only,
                                   torch.from_numpy(a), torch.from_numpy(b))
        Xs.append(X); ys.append(y); fs.append(f); ps.append(p)
    return torch.cat(Xs), torch.cat(ys), torch.cat(fs), torch.cat(ps)


@dataclass
class StepRecord:
    losses: dict
    grad_norm: float | None


def dropout_mask(seed, layer, n_seq, T, hidden, p):
    """Keep-mask of the input of layer `layer` (>=1), shape (T, N, H), values 0 or 1/(1-p), exactly as the
    kernels regenerate it: Philox counter (row_lo, row_hi, t | layer<<16, unit//8), unit j of the block keeps
    iff its 16-bit lane >= round(p*65536).  Integer work: compared bit-exactly with b200_dropout_mask."""
    rows = np.arange(n_seq, dtype=np.uint64)[None, :, None]
    ts = np.arange(T, dtype=np.uint32)[:, None, None]
    ub = np.arange(hidden // 8, dtype=np.uint32)[None, None, :]
    k0, k1 = seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF
    x = philox4x32_10((rows & np.uint64(0xFFFFFFFF)).astype(np.uint32), (rows >> np.uint64(32)).astype(np.uint32),
                      ts | np.uint32(layer << 16), ub, k0, k1)
    lanes = []
    for w in x:
        lanes += [w & np.uint32(0xFFFF), w >> np.uint32(16)]
    u16 = np.stack(lanes, axis=-1).reshape(T, n_seq, hidden)
    thr = min(65535, int(np.float32(p) * np.float32(65536.0) + np.float32(0.5)))
    return np.where(u16 >= thr, np.float32(1.0 / (1.0 - p)), np.float32(0.0)).astype(np.float32)
