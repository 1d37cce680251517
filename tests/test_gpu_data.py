"""-m gpu: file-backed / series batches, five features and the reconstruction split (f4) through the
C-ABI (b200_series_batch, b200_gen_series) against golden vectors made by the reference's own prepare_data body
(oracle/make_golden.py::variants) and against the oracle."""
import numpy as np
import pytest
import torch

import lstm_oracle as O
import masters_thesis_b200 as B
from masters_thesis_b200._lib import check, load, ptr
from _gpu_util import DEV, oracle_model, rel_err

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def variants(golden_dir):
    g = np.load(golden_dir / "golden_variants.npz")
    return {k: torch.from_numpy(g[k]) for k in g.files}


@pytest.mark.parametrize("tag", ["f5", "recon", "real"])
def test_series_batch_matches_reference_golden(tag, variants):
    T, n, stride, pred, inter = [int(v) for v in variants[f"{tag}_cfg"]]
    src = B.data.SeriesBatchSource(DEV, variants["r_stocks"], variants["r_market"], T, n, stride, 3 if inter else 5, bool(pred))
    Xg, yg = variants[f"{tag}_X"], variants[f"{tag}_y"]
    assert src.n_windows == Xg.shape[0]
    # whole range, then an arbitrary index list (ragged: 3 windows, out of order)
    for idx in (None, torch.tensor([Xg.shape[0] - 1, 0, 2])):
        b = src.generate(first_window=0, n_windows=Xg.shape[0]) if idx is None else src.generate(window_idx=idx)
        sel = slice(None) if idx is None else idx
        X, y, factor, inv_psi = [t.cpu() for t in b.to_reference()]
        assert torch.equal(X, Xg[sel])                                   # byte copies and one fp32 product per feature
        assert torch.equal(y[..., :2], yg[sel][..., :2])
        nB, S = X.shape[:2]
        np.testing.assert_allclose(b.ols_alpha.cpu().reshape(nB, S), yg[sel][:, :, 0, 2], rtol=2e-4, atol=2e-6)
        np.testing.assert_allclose(b.ols_beta.cpu().reshape(nB, S), yg[sel][:, :, 0, 3], rtol=2e-4, atol=2e-6)
        np.testing.assert_allclose(factor, variants[f"{tag}_factor"][sel], rtol=1e-5, atol=1e-7)
        np.testing.assert_allclose(inv_psi, variants[f"{tag}_inv_psi"][sel], rtol=2e-4)


def test_series_batch_rejects_bad_arguments(variants):
    src = B.data.SeriesBatchSource(DEV, variants["r_stocks"], variants["r_market"], 60, 30, 45, 3, True)
    with pytest.raises(B.B200Error):
        src.generate(first_window=src.n_windows - 1, n_windows=2)        # runs past the end of the series
    with pytest.raises(IndexError):
        src.generate(window_idx=torch.tensor([src.n_windows]))
    with pytest.raises(ValueError):
        B.data.SeriesBatchSource(DEV, variants["r_stocks"], variants["r_market"][:-1])
    empty = src.generate(first_window=0, n_windows=0)                     # empty batch is a no-op
    assert empty.x.shape[1] == 0


def test_gen_series_is_the_series_behind_gen_batch():
    """b200_gen_series(day range) and b200_gen_batch(windows) address the same Philox series: bit-identical."""
    S, T, n, stride, seed = 100, 60, 30, 90, 4321
    gsrc = B.data.DeviceBatchSource(DEV, seed, S, T, n, stride, 3)
    stocks, market = B.data.generator_series(DEV, seed, S, 5 * stride + 7)
    ssrc = B.data.SeriesBatchSource(DEV, stocks, market, T, n, stride, 3, True)
    a, b = gsrc.generate(first_window=0, n_windows=5), ssrc.generate(first_window=0, n_windows=5)
    for name in ("x", "y_stock", "y_market", "factor", "inv_psi"):
        assert torch.equal(getattr(a, name), getattr(b, name)), name
    # a later day range starts mid-series
    st2, mk2 = B.data.generator_series(DEV, seed, S, 90, first_day=2 * stride)
    assert torch.equal(st2, stocks[:, 2 * stride:2 * stride + 90]) and torch.equal(mk2, market[2 * stride:2 * stride + 90])


@pytest.mark.parametrize("engine", ["ffma", "tc_precise"])
def test_training_step_on_real_layout_matches_oracle(engine, variants):
    """datamodule=real shape (25 portfolios, stride 5): one combined-objective step, gradients vs the oracle."""
    if engine != "ffma" and not load().b200_mode_available(B.MODES[engine]):
        pytest.skip("tcgen05 engine unavailable")
    T, n, stride = 60, 30, 5
    src = B.data.SeriesBatchSource(DEV, variants["r_stocks"], variants["r_market"], T, n, stride, 3, True)
    idx = torch.tensor([5, 17, 40])
    b = src.generate(window_idx=idx)
    torch.manual_seed(3)
    model = B.FinancialLstmCombined(3, 64, 2, 0.0, 1e-4, mse_weight=100)
    init = model.flat_parameters().clone()
    e = model.attach(DEV, engine, T, n, 25)
    e.forward(b, True, 0)
    losses = e.heads(b, "combined", 100.0, None, 1.0, want_grad=True)
    e.backward(b, True, 0)
    om = oracle_model(init, 2, 0.0)
    ref = [variants["real_X"][idx], variants["real_y"][idx], variants["real_factor"][idx], variants["real_inv_psi"][idx]]
    ls = O.batch_losses(om, ref, "combined", 100.0)
    ls["total"].backward()
    assert float(losses[2]) == pytest.approx(float(ls["total"].detach()), rel=1e-4)
    assert rel_err(e.grads[: e.P].cpu().numpy(), O.flat_grads(om).numpy()) < 1e-3


def test_datamodule_from_files_and_reconstruction(tmp_path, variants):
    """The reference's on-disk layout (data_dir/stocks.pt, market.pt) drives the same loaders; prediction_task=false
    with the generator materialises its series first."""
    d = tmp_path / "real"
    d.mkdir()
    torch.save(variants["r_stocks"], d / "stocks.pt")
    torch.save(variants["r_market"], d / "market.pt")
    dm = B.FinancialLstmDataModule(d, 60, 30, 5, True, True, batch_size=4, device=DEV)
    dm.prepare_data(); dm.setup("fit")
    assert (dm.source_kind, dm.n_stocks, dm.n_windows) == ("files", 25, 63)
    assert (dm.train_end, dm.val_end) == (44, 56)
    nb = sum(1 for _ in dm.train_dataloader())
    assert nb == 11
    model = B.FinancialLstmNll(3, 64, 2, 0.0, 1e-4)
    tr = B.B200Trainer(max_epochs=1, gradient_clip_val=5.0, engine="ffma", verbose=False)
    tr.fit(model, dm)
    tr.test(model, dm)
    assert np.isfinite(float(tr.checkpoint_callback.best_model_score)) and np.isfinite(model.test_results["test_loss"]["nll"])

    rec = B.FinancialLstmDataModule(tmp_path / "synthetic", 60, 30, 60, False, True, batch_size=2, seed=5, n_samples=600,
                                    device=DEV)
    rec.prepare_data(); rec.setup("fit")
    assert rec.source_kind == "generator-series" and rec.n_windows == 10
    b = next(iter(rec.val_dataloader()))
    X, y, _, _ = [t.cpu() for t in b.to_reference()]
    assert torch.equal(y[0, :, :, 0], X[0, :, 30:, 0])      # the target IS the tail of the look-back


@pytest.mark.parametrize("B_,S,T,n,F", [(3, 7, 60, 30, 3), (2, 100, 60, 30, 5), (1, 33, 45, 10, 3)])
def test_collate_reference_kernel(B_, S, T, n, F):
    """b200_collate_reference (device layout change of a reference-format batch) against the torch permute it replaces."""
    g = torch.Generator().manual_seed(B_ * 1000 + S)
    X, y = torch.randn(B_, S, T, F, generator=g), torch.randn(B_, S, n, 4, generator=g)
    y[:, :, :, 1] = y[:, :1, :, 1]                      # the market channel is the same for every stock of a window
    factor, inv_psi = torch.randn(B_, 2, generator=g), torch.rand(B_, S, generator=g)
    b = B.DeviceBatch.from_reference([X, y, factor, inv_psi], DEV)
    assert torch.equal(b.x.cpu(), X.reshape(B_ * S, T, F).permute(1, 0, 2).contiguous())
    assert torch.equal(b.y_stock.cpu(), y[..., 0].reshape(B_ * S, n))
    assert torch.equal(b.y_market.cpu(), y[:, 0, :, 1])
    assert torch.equal(b.inv_psi.cpu(), inv_psi.reshape(-1)) and torch.equal(b.factor.cpu(), factor)
    back = b.to_reference()
    assert torch.equal(back[0].cpu(), X) and torch.equal(back[1].cpu()[..., :2], y[..., :2])
