"""-m gpu: the training step at BASELINE.json's full size (configs[1]: 4096 windows = 409,600 sequences x 60 steps per
GPU, 25 GB of stash, element offsets beyond 2^31) through size-independent properties: rows are independent in the
forward pass (spot rows vs the oracle), the mse gradient is linear in the batch (two halves vs the whole), and
gradients are bit-identical run to run."""
import numpy as np
import pytest
import torch

import lstm_oracle as O
import masters_thesis_b200 as B
from masters_thesis_b200._lib import load
from _gpu_util import DEV, oracle_model, rel_err

pytestmark = pytest.mark.gpu
BW, S, SEED = 4096, 100, 1234


@pytest.mark.parametrize("engine", ["tc_precise"])
def test_full_size_step_properties(engine):
    if not load().b200_mode_available(B.MODES[engine]):
        pytest.skip("tcgen05 engine unavailable")
    free, _ = torch.cuda.mem_get_info(0)
    if free < 110e9:
        pytest.skip("needs ~95 GB of device memory")
    torch.manual_seed(7)
    model = B.FinancialLstmMse(3, 64, 2, 0.0, 1e-4)
    flat = model.flat_parameters().clone()
    e = model.attach(DEV, engine)
    src = B.data.DeviceBatchSource(DEV, SEED)
    full = src.generate(first_window=0, n_windows=BW)

    def step(batch, scale):
        e.forward(batch, training=True, dropout_seed=1)
        losses = e.heads(batch, "mse", 100.0, None, scale, True).clone()
        e.backward(batch, True, 1)
        return losses, e.grads[: e.P].clone()

    l_full, g_full = step(full, 1.0)
    # (1) forward rows are independent: windows at the start, in the middle and at the very end against the oracle on
    #     the same windows
    N = BW * S
    w_mid = 1234                                # every layer-1 stash offset at t >= 22 is beyond 2^31 elements at this size
    spots = [0, 1, w_mid, BW // 2, BW - 1]
    a_dev, b_dev = e.alpha[:N].cpu(), e.beta[:N].cpu()
    om = oracle_model(flat, 2, 0.0)
    X, _, _, _ = O.generate_windows(spots, seed=SEED)
    with torch.no_grad():
        om.eval()
        for k, w in enumerate(spots):
            ao, bo = om(X[k])
            np.testing.assert_allclose(a_dev[w * S:(w + 1) * S].numpy(), ao[:, 0].numpy(), atol=2e-5, err_msg=f"alpha window {w}")
            np.testing.assert_allclose(b_dev[w * S:(w + 1) * S].numpy(), bo[:, 0].numpy(), atol=2e-5, err_msg=f"beta window {w}")
    # (2) determinism at full size
    l2, g2 = step(full, 1.0)
    assert torch.equal(g_full, g2) and torch.equal(l_full, l2)
    # (3) linearity in the batch: gradient(whole) = 0.5 gradient(first half) + 0.5 gradient(second half); same for the loss
    del full
    h0 = src.generate(first_window=0, n_windows=BW // 2)
    l_a, g_a = step(h0, 0.5)
    h1 = src.generate(first_window=BW // 2, n_windows=BW // 2)
    l_b, g_b = step(h1, 0.5)
    assert rel_err((g_a + g_b).cpu().numpy(), g_full.cpu().numpy()) < 2e-5
    assert float(l_full[0]) == pytest.approx(0.5 * (float(l_a[0]) + float(l_b[0])), rel=1e-6)
    assert np.isfinite(g_full.cpu().numpy()).all() and float(g_full.abs().max()) > 0


def test_full_size_fixed_budget_vs_oracle():
    """The headline configuration under the north_star gate, on the engine and through the object the bench uses
    (BASELINE.json configs[1]: small model, mse, dropout 0.2, 4096 windows = 409,600 sequences per step, tc_precise,
    B200Trainer): five optimisation steps against the CPU oracle accumulated over 32-window chunks (fp64 sums across chunks) on the same windows and
    the same Philox dropout masks -- per-step loss within 1e-4 relative, the FULL step-0 gradient within 1e-3 (whole
    vector and per tensor), then best_val_loss, test mae and test nll (train.py:204-211) and the final parameters."""
    if not load().b200_mode_available(B.MODES["tc_precise"]):
        pytest.skip("tcgen05 engine unavailable")
    free, _ = torch.cuda.mem_get_info(0)
    if free < 120e9:
        pytest.skip("needs ~105 GB of device memory")
    from _budget import check_budget, run_budget
    rec = run_budget(DEV, BW, 5, 0.2, engine="tc_precise", kind="mse", want_grads_at=(0,))
    check_budget(rec)
