"""-m gpu: SURVEY.md f1 -- the checkpoint and TensorBoard output of `train.py trainer=b200` and their consumer, the
evaluation script `test.py`, against the oracle restatement of the reference script (oracle.script_test_windows)."""
import numpy as np
import pytest
import torch

import lstm_oracle as O
import masters_thesis_b200 as B
from _gpu_util import DEV, oracle_model

pytestmark = pytest.mark.gpu


def _scalars(log_dir):
    from tensorboard.backend.event_processing.event_accumulator import EventAccumulator
    acc = EventAccumulator(str(log_dir), size_guidance={"scalars": 0, "histograms": 0})
    acc.Reload()
    return acc


def test_train_then_test_script_roundtrip(tmp_path, monkeypatch):
    """train.py (2 epochs, tiny series) -> checkpoint + event file -> test.py on that checkpoint -> the per-window
    evaluation quantities equal the oracle's on the same windows and weights."""
    import train as train_script
    import test as test_script
    monkeypatch.chdir(tmp_path)
    ov = ["trainer=b200", "trainer.max_epochs=2", "trainer.engine=ffma", "loss=combined", "datamodule.batch_size=2",
          f"logger.save_dir={tmp_path / 'logs'}", "trainer.n_samples=3600", "trainer.seed=77"]
    best = train_script.main(ov)
    run_dir = tmp_path / "logs" / "FinancialLstm" / "synthetic" / "combined_small_lr0.0001_b200"
    cks = sorted((run_dir / "checkpoints").glob("*.ckpt"))
    assert any(c.name == "last.ckpt" for c in cks) and any(c.name.startswith("best_epoch=") for c in cks)
    ck = torch.load(run_dir / "checkpoints" / "last.ckpt", weights_only=False)
    assert list(ck["state_dict"].keys()) == list(O.OracleLstm(3, 64, 2, 0.2).state_dict().keys())     # Lightning layout
    assert ck["hyper_parameters"]["hidden_size"] == 64 and ck["hyper_parameters"]["mse_weight"] == 100
    acc = _scalars(run_dir)
    tags = set(acc.Tags()["scalars"])
    assert {"loss/total/train", "loss/total/val", "loss/mse/val", "loss/nll/val", "lr-Adam"} <= tags    # train.py:150-166
    assert len(acc.Scalars("loss/total/val")) == 2
    best_ck = [c for c in cks if c.name.startswith("best_epoch=")][-1]
    assert f"{best:.4f}" in best_ck.name                                                               # filename template, train.py:153

    summary = test_script.main(ov + [f"checkpoint={best_ck}"])
    res = np.load(run_dir / "test_results.npz")
    # oracle on the same test windows with the checkpoint's weights
    model = B.FinancialLstmCombined.load_from_checkpoint(best_ck)
    om = oracle_model(model.flat_parameters(), 2, 0.0)
    n_win = (3600 - 90) // 90 + 1
    X, y, _, _ = O.generate_windows(list(range(int(0.9 * n_win), n_win)), seed=77)
    ref = O.script_test_windows(om, X, y)
    for k, sub in ref.items():
        for s, v in sub.items():
            np.testing.assert_allclose(res[f"{k}__{s}"], v.numpy(), rtol=2e-3, atol=2e-5, err_msg=f"{k}/{s}")
    assert summary["estimation/beta/ols_rmse"] == pytest.approx(float(((ref["beta"]["ols"] - ref["beta"]["true"]) ** 2).mean().sqrt()), rel=1e-3)
    assert "hist/betas/model" in set(_scalars(run_dir).Tags()["histograms"])


def test_reference_style_checkpoint_loads(tmp_path):
    """A checkpoint in the reference's (Lightning) layout -- state_dict of the torch modules + hyper_parameters --
    loads into the mirror and evaluates identically."""
    torch.manual_seed(11)
    om = O.OracleLstm(3, 64, 2, 0.2)
    path = tmp_path / "ref.ckpt"
    torch.save({"state_dict": om.state_dict(), "hyper_parameters": {"input_size": 3, "hidden_size": 64, "num_layers": 2,
                                                                     "dropout": 0.2, "learning_rate": 1e-4, "weight_decay": 1e-5},
                "epoch": 3, "global_step": 12, "pytorch-lightning_version": "2.6.0"}, path)
    model = B.FinancialLstmMse.load_from_checkpoint(path, map_location="cpu")
    model.attach(DEV, "ffma")
    x = torch.randn(50, 60, 3)
    a, b = model(x)
    om.eval()
    with torch.no_grad():
        ao, bo = om(x)
    np.testing.assert_allclose(a.cpu().numpy(), ao.numpy(), atol=1e-5)
    np.testing.assert_allclose(b.cpu().numpy(), bo.numpy(), atol=1e-5)


def test_loss_mailbox_matches_direct_reads():
    """The pipelined per-step loss read-back returns every step's values, in order, equal to direct reads."""
    torch.manual_seed(3)
    model = B.FinancialLstmMse(3, 64, 2, 0.0, 1e-4)
    dm = B.FinancialLstmDataModule(lookback_window=60, target_window=30, stride=90, batch_size=2, seed=5, n_samples=90 * 20,
                                   device=DEV)
    tr = B.B200Trainer(max_epochs=1, gradient_clip_val=5.0, engine="ffma", verbose=False)
    dm.prepare_data(); dm.setup("fit")
    tr._attach(model, dm)
    box, got, want = B.LossMailbox(DEV, depth=2), [], []
    for k in range(5):
        b = dm.source.generate(first_window=2 * k, n_windows=2)
        losses = tr.training_step(model, b, 1e-4)
        want.append(losses.clone())
        got += box.post(losses)
    got += box.drain()
    assert len(got) == 5
    for g, w in zip(got, want):
        assert g == w.cpu().tolist()


def test_warm_start_from_synthetic_checkpoint_on_a_real_layout_series(tmp_path, monkeypatch):
    """SURVEY.md f3, the thesis' warm-up recipe (tex/diplomski_rad.tex:1134-1147): stage 1 trains on the synthetic datamodule,
    stage 2 continues on `datamodule=real` (here a 25-portfolio series saved in the reference's stocks.pt / market.pt
    layout, stride 5) from stage 1's best checkpoint given as `checkpoint=`.  The second run must START from the first
    run's parameters: with a zero learning rate its checkpoint equals stage 1's."""
    import train as train_script
    monkeypatch.chdir(tmp_path)
    common = ["trainer=b200", "trainer.engine=ffma", "trainer.max_epochs=1", f"logger.save_dir={tmp_path / 'logs'}", "trainer.seed=5"]
    train_script.main(common + ["loss=combined", "datamodule.batch_size=4", "trainer.n_samples=1800"])
    run1 = tmp_path / "logs" / "FinancialLstm" / "synthetic" / "combined_small_lr0.0001_b200"
    best = sorted((run1 / "checkpoints").glob("best_*.ckpt"))[-1]
    real = tmp_path / "data" / "real"
    real.mkdir(parents=True)
    stocks, market = B.data.generator_series(DEV, 9, 25, 700)
    torch.save(stocks.cpu(), real / "stocks.pt")
    torch.save(market.cpu(), real / "market.pt")
    train_script.main(common + ["datamodule=real", f"datamodule.data_dir={real}", "loss=mse", "datamodule.batch_size=8",
                                "model.learning_rate=0.0", f"checkpoint={best}"])
    run2 = tmp_path / "logs" / "FinancialLstm" / "real" / "mse_small_lr0.0_b200"
    sd1 = torch.load(best, weights_only=False)["state_dict"]
    sd2 = torch.load(run2 / "checkpoints" / "last.ckpt", weights_only=False)["state_dict"]
    assert list(sd1) == list(sd2)
    for k in sd1:
        assert torch.equal(sd1[k], sd2[k]), k
