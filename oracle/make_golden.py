"""Generate tests/golden/*.npz by running the REFERENCE's own modules (imported unmodified
from /root/reference behind oracle/ref_stubs.py).  Runs only in the build container.

    python oracle/make_golden.py

What is recorded (all fp32 CPU, torch.manual_seed set explicitly because the reference
never seeds; float32 matmul precision 'highest'):

  golden_data.npz    a 6-window synthetic series drawn by SyntheticLogReturns.generate and
                     pushed through lookback_target_split / add_quadratic_features /
                     ols_features / _append_feature (the prepare_data body, src/data.py:193-214)
  golden_<loss>.npz  FinancialLstm{Mse,Nll,Combined} small model (dropout 0), init state,
                     3 training steps at batch_size 2 in Lightning's step order with
                     gradient_clip_val 5.0 (trainer=fast): per-step losses, step-0 gradients
                     (pre-clip) and norm, parameters after 3 steps, validation means over
                     windows 4,5 and test_step metrics over windows 4,5
"""
from __future__ import annotations

import sys
from pathlib import Path

import numpy as np
import torch

HERE = Path(__file__).resolve().parent
sys.path.insert(0, str(HERE))
import ref_stubs  # noqa: E402

OUT = HERE.parent / "tests" / "golden"
SEED_DATA, SEED_MODEL = 20260101, 1234
N_WIN, BATCH, STEPS, CLIP = 6, 2, 3, 5.0


def main():
    src = ref_stubs.install()
    from src.common import lookback_target_split, add_quadratic_features, ols_features
    from src.data import _append_feature

    torch.set_num_threads(1)
    torch.set_float32_matmul_precision("highest")
    OUT.mkdir(parents=True, exist_ok=True)

    torch.manual_seed(SEED_DATA)
    r_stocks, r_market, alphas, betas = src.SyntheticLogReturns.generate(100, 90 * N_WIN)
    X, y = lookback_target_split(r_stocks, r_market, 60, 30, 90, True)
    X = add_quadratic_features(X, interaction_only=True)
    a_hat, b_hat, factor, inv_psi = ols_features(y)
    y = _append_feature(y, alphas)
    y = _append_feature(y, betas)
    np.savez_compressed(OUT / "golden_data.npz",
                        r_stocks=r_stocks.numpy(), r_market=r_market.numpy(),
                        alphas=alphas.numpy(), betas=betas.numpy(),
                        X=X.numpy(), y=y.numpy(), factor=factor.numpy(), inv_psi=inv_psi.numpy(),
                        ols_alpha=a_hat.numpy(), ols_beta=b_hat.numpy())

    classes = {"mse": src.FinancialLstmMse, "nll": src.FinancialLstmNll,
               "combined": src.FinancialLstmCombined}
    order = [(0, 2), (2, 4), (0, 2)]
    for kind, cls in classes.items():
        torch.manual_seed(SEED_MODEL)
        model = cls(input_size=3, hidden_size=64, num_layers=2, dropout=0.0, learning_rate=1e-4)
        names = [n for n, _ in model.named_parameters()]
        init = torch.cat([p.detach().reshape(-1) for p in model.parameters()]).numpy().copy()
        opt = model.configure_optimizers()["optimizer"]
        model.train()
        rec = {"losses": [], "mse": [], "nll": []}
        grads0 = gnorm0 = None
        for s in range(STEPS):
            lo, hi = order[s]
            batch = [X[lo:hi], y[lo:hi], factor[lo:hi], inv_psi[lo:hi]]
            loss = model.training_step(batch, s)
            opt.zero_grad()
            loss.backward()
            if s == 0:
                grads0 = torch.cat([p.grad.reshape(-1) for p in model.parameters()]).numpy().copy()
            gn = torch.nn.utils.clip_grad_norm_(model.parameters(), CLIP)
            if s == 0:
                gnorm0 = float(gn)
            opt.step()
            rec["losses"].append(float(loss.detach()))
            for k in ("mse", "nll"):
                m = model.logged.get(f"loss/{k}/train")
                rec[k].append(float(m.compute()) if m is not None else np.nan)
        final = torch.cat([p.detach().reshape(-1) for p in model.parameters()]).numpy().copy()

        # validation (batch 1, eval) over windows 4,5 -> epoch means as Lightning logs them
        model.eval()
        val_total = []
        with torch.no_grad():
            for w in (4, 5):
                batch = [X[w:w + 1], y[w:w + 1], factor[w:w + 1], inv_psi[w:w + 1]]
                model.logged.clear()
                model.validation_step(batch, w)
                val_total.append(float(model.logged["loss/total/val"]))
            model.on_test_start()
            for w in (4, 5):
                batch = [X[w:w + 1], y[w:w + 1], factor[w:w + 1], inv_psi[w:w + 1]]
                model.test_step(batch, w)
            model.on_test_end()
            a_out, b_out = model(X[4])
        np.savez_compressed(
            OUT / f"golden_{kind}.npz",
            names=np.array(names), init=init, grads0=grads0, gnorm0=np.float64(gnorm0),
            final=final, losses=np.array(rec["losses"]), mse=np.array(rec["mse"]), nll=np.array(rec["nll"]),
            val_total=np.array(val_total),
            test_mae=np.float64(model.test_results["test_loss"]["mae"]),
            test_nll=np.float64(model.test_results["test_loss"]["nll"]),
            alpha_w4=a_out.numpy(), beta_w4=b_out.numpy(),
            meta=np.array([SEED_DATA, SEED_MODEL, N_WIN, BATCH, STEPS]), clip=np.float64(CLIP))
        print(kind, rec["losses"], "gnorm0", gnorm0, "val", val_total,
              "test", model.test_results["test_loss"])


def variants():
    """golden_variants.npz: the reference's prepare_data body (src/data.py:193-214) on a 25-stock, 400-day slice of the
    golden series for the switches no shipped config uses: all five quadratic features,
    the reconstruction split (prediction=False), and no true coefficients (real data: y channels 2, 3 = per-window OLS)."""
    ref_stubs.install()
    from src.common import lookback_target_split, add_quadratic_features, ols_features
    from src.data import _append_feature
    torch.set_num_threads(1)
    torch.set_float32_matmul_precision("highest")
    d = np.load(OUT / "golden_data.npz")
    rs, rm = torch.from_numpy(d["r_stocks"])[:25, :400].contiguous(), torch.from_numpy(d["r_market"])[:400].contiguous()
    out = {"r_stocks": rs.numpy(), "r_market": rm.numpy()}
    for tag, (T, n, stride, pred, inter) in {"f5": (60, 30, 45, True, False), "recon": (60, 30, 45, False, True),
                                             "real": (60, 30, 5, True, True)}.items():
        X, y = lookback_target_split(rs, rm, T, n, stride, pred)
        X = add_quadratic_features(X, interaction_only=inter)
        a_hat, b_hat, factor, inv_psi = ols_features(y)
        y = _append_feature(_append_feature(y, a_hat), b_hat)
        out.update({f"{tag}_X": X.numpy(), f"{tag}_y": y.numpy(), f"{tag}_factor": factor.numpy(),
                    f"{tag}_inv_psi": inv_psi.numpy(), f"{tag}_cfg": np.array([T, n, stride, int(pred), int(inter)])})
        print(tag, tuple(X.shape), tuple(y.shape))
    np.savez_compressed(OUT / "golden_variants.npz", **out)


BATCH_CASES = [(kind, 64, 2, B) for kind in ("mse", "nll", "combined") for B in (1, 8, 64)] + [("combined", 128, 3, 8)]
BATCH_SEED_DATA, BATCH_FIRST_WINDOW = 4321, 500


def batch_goldens():
    """golden_batches.npz: the reference's FinancialLstm{Mse,Nll,Combined}, dropout 0, for
    batch_size B in {1, 8, 64} on the small model plus one widened shape (hidden 128 x 3 layers): 3 optimisation steps
    in Lightning's order with gradient_clip_val 5.0 -- per-step losses, step-0 gradients (pre-clip) and their norm.
    Inputs are NOT stored: step s uses windows BATCH_FIRST_WINDOW + s*B .. + B of the written generator specification
    (oracle/lstm_oracle.py generate_days, seed BATCH_SEED_DATA), pushed here through the REFERENCE's own
    lookback_target_split / add_quadratic_features / ols_features / _append_feature; the tests rebuild the same
    batches with lstm_oracle.generate_windows (pinned bit-equal to the reference's prep by test_oracle.py)."""
    src = ref_stubs.install()
    from src.common import lookback_target_split, add_quadratic_features, ols_features
    from src.data import _append_feature
    import lstm_oracle as O
    torch.set_num_threads(8)
    torch.set_float32_matmul_precision("highest")
    classes = {"mse": src.FinancialLstmMse, "nll": src.FinancialLstmNll, "combined": src.FinancialLstmCombined}

    def ref_batch(first, B):
        r_s, r_m, a, b = O.generate_days(first * 90, 90 * B, 100, BATCH_SEED_DATA)
        X, y = lookback_target_split(torch.from_numpy(r_s), torch.from_numpy(r_m), 60, 30, 90, True)
        X = add_quadratic_features(X, interaction_only=True)
        _, _, factor, inv_psi = ols_features(y)
        y = _append_feature(_append_feature(y, torch.from_numpy(a)), torch.from_numpy(b))
        return [X.contiguous(), y.contiguous(), factor, inv_psi]

    out = {"cases": np.array([f"{k}_h{H}_l{L}_b{B}" for k, H, L, B in BATCH_CASES]),
           "meta": np.array([BATCH_SEED_DATA, SEED_MODEL, BATCH_FIRST_WINDOW, STEPS]), "clip": np.float64(CLIP)}
    for kind, H, L, B in BATCH_CASES:
        tag = f"{kind}_h{H}_l{L}_b{B}"
        torch.manual_seed(SEED_MODEL)
        model = classes[kind](input_size=3, hidden_size=H, num_layers=L, dropout=0.0, learning_rate=1e-4)
        opt = model.configure_optimizers()["optimizer"]
        model.train()
        losses, grads0, gnorm0 = [], None, None
        for s in range(STEPS):
            batch = ref_batch(BATCH_FIRST_WINDOW + s * B, B)
            loss = model.training_step(batch, s)
            opt.zero_grad()
            loss.backward()
            if s == 0:
                grads0 = torch.cat([p.grad.reshape(-1) for p in model.parameters()]).numpy().copy()
            gn = torch.nn.utils.clip_grad_norm_(model.parameters(), CLIP)
            if s == 0:
                gnorm0 = float(gn)
            opt.step()
            losses.append(float(loss.detach()))
        out[f"{tag}_losses"] = np.array(losses)
        out[f"{tag}_grads0"] = grads0
        out[f"{tag}_gnorm0"] = np.float64(gnorm0)
        print(tag, losses, gnorm0, flush=True)
    np.savez_compressed(OUT / "golden_batches.npz", **out)


if __name__ == "__main__":
    if "--variants-only" in sys.argv:
        variants()
    elif "--batches-only" in sys.argv:
        batch_goldens()
    else:
        main()
        variants()
        batch_goldens()
