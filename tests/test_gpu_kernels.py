"""-m gpu: every bandwidth-bound kernel against the oracle, through the C-ABI."""
import numpy as np
import pytest
import torch

import lstm_oracle as O
import masters_thesis_b200 as B
from masters_thesis_b200._lib import check, load, ptr
from _gpu_util import DEV, rel_err

pytestmark = pytest.mark.gpu
torch.set_float32_matmul_precision("highest")


def test_philox_bit_exact():
    lib = load()
    rng = np.random.default_rng(0)
    ctr = rng.integers(0, 2**32, size=(4096, 4), dtype=np.uint64).astype(np.uint32)
    ctr[0] = 0
    ctr[1] = 0xFFFFFFFF
    seed = 0x0123456789ABCDEF
    d_ctr = torch.from_numpy(ctr.view(np.int32)).to(DEV)
    d_out = torch.empty_like(d_ctr)
    check(lib.b200_philox4x32(seed, ptr(d_ctr), ctr.shape[0], ptr(d_out), None), "philox")
    got = d_out.cpu().numpy().view(np.uint32)
    want = np.stack(O.philox4x32_10(ctr[:, 0], ctr[:, 1], ctr[:, 2], ctr[:, 3], seed & 0xFFFFFFFF, seed >> 32), axis=1)
    assert np.array_equal(got, want)


@pytest.mark.parametrize("F", [3, 5])
def test_gen_batch_matches_spec(F):
    src = B.data.DeviceBatchSource(DEV, seed=1234, n_features=F)
    idx = torch.tensor([0, 3, 11110, 7, 2**33 + 5])
    b = src.generate(window_idx=idx)
    torch.cuda.synchronize()
    X, y, factor, inv_psi = O.generate_windows(idx.tolist(), seed=1234, interaction_only=(F == 3))
    a, be = src.coefficients()
    got = b.to_reference(a, be)
    a_ref, b_ref = O.stock_coefficients(100, 1234)
    np.testing.assert_allclose(a.cpu().numpy(), a_ref, rtol=0, atol=2e-6)
    np.testing.assert_allclose(be.cpu().numpy(), b_ref, rtol=0, atol=2e-6)
    # float transform (log/sincos/sqrt) differs by a few ulp between libm and CUDA; heavy tails -> abs+rel
    np.testing.assert_allclose(got[0].cpu().numpy(), X.numpy(), rtol=2e-4, atol=2e-5)
    np.testing.assert_allclose(got[1].cpu().numpy(), y.numpy(), rtol=2e-4, atol=2e-5)
    np.testing.assert_allclose(got[2].cpu().numpy(), factor.numpy(), rtol=1e-3, atol=1e-5)
    np.testing.assert_allclose(got[3].cpu().numpy(), inv_psi.numpy(), rtol=2e-3)
    # OLS targets against the reference restatement evaluated on the DEVICE's own windows (same inputs)
    _, _, f2, p2 = O.ols_targets(got[1].cpu()[..., :2])
    np.testing.assert_allclose(got[2].cpu().numpy(), f2.numpy(), rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(got[3].cpu().numpy(), p2.numpy(), rtol=2e-4)


def test_gen_batch_contiguous_equals_indexed_and_edge_sizes():
    src = B.data.DeviceBatchSource(DEV, seed=99)
    a = src.generate(first_window=5, n_windows=3)
    ax = [t.clone() for t in (a.x, a.y_stock, a.y_market, a.factor, a.inv_psi)]
    b = src.generate(window_idx=torch.tensor([5, 6, 7]))
    for u, v in zip(ax, (b.x, b.y_stock, b.y_market, b.factor, b.inv_psi)):
        assert torch.equal(u, v)
    one = src.generate(first_window=6, n_windows=1)
    assert torch.equal(one.y_stock, ax[1][100:200])
    # invalid arguments are rejected with a message, not a crash
    lib = load()
    rc = lib.b200_gen_batch(1, None, 0, 1, 100, 60, 30, 90, 4, ptr(a.x), ptr(a.y_stock), ptr(a.y_market), ptr(a.factor), ptr(a.inv_psi), None)
    assert rc == -1 and b"n_features" in lib.b200_last_error()


def test_dropout_mask_bit_exact():
    lib = load()
    N, T, H, p, seed, layer = 37, 5, 64, 0.2, 0xDEADBEEF12345, 3
    out = torch.empty(T, N, H, device=DEV)
    check(lib.b200_dropout_mask(seed, layer, N, T, H, p, ptr(out), None), "dropout_mask")
    want = O.dropout_mask(seed, layer, N, T, H, p)
    assert np.array_equal(out.cpu().numpy(), want)
    keep = (want > 0).mean()
    assert abs(keep - 0.8) < 0.02


def _heads_case(B_, S, n, H, seed):
    g = torch.Generator().manual_seed(seed)
    hT = torch.tanh(torch.randn(B_ * S, H, generator=g))
    hp = torch.randn(2 * H + 2, generator=g) * 0.2
    y_stock = torch.randn(B_ * S, n, generator=g) * 0.6
    y_market = torch.randn(B_, n, generator=g) * 0.5 + 0.07
    factor = torch.stack([y_market.mean(1), y_market.var(1)], 1)
    inv_psi = 1.0 / (0.05 + torch.rand(B_ * S, generator=g))
    return hT, hp, y_stock, y_market, factor, inv_psi


def _heads_oracle(kind, w, hT, hp, y_stock, y_market, factor, inv_psi, B_, S, wpg):
    H = hT.shape[1]
    hT = hT.double().requires_grad_()
    hp = hp.double().requires_grad_()
    alpha = hT @ hp[:H] + hp[H]
    beta = hT @ hp[H + 1:2 * H + 1] + hp[2 * H + 1]
    m_rows = y_market.double().repeat_interleave(S, 0)
    mse = torch.mean((alpha[:, None] + beta[:, None] * m_rows - y_stock.double()) ** 2)
    mae = torch.mean(torch.abs(alpha[:, None] + beta[:, None] * m_rows - y_stock.double()))
    G = B_ // wpg
    nll = 0
    for g in range(G):
        r = slice(g * wpg * S, (g + 1) * wpg * S)
        fm, fv = factor[g * wpg, 0].double(), factor[g * wpg, 1].double()
        inv_cov = O.woodbury_inverse_cov(beta[r, None], torch.diag(inv_psi[r].double()), fv)
        nll = nll + O.gaussian_nll_dense((alpha[r] + beta[r] * fm)[:, None], inv_cov, y_stock[r].double())
    nll = nll / G
    total = {"mse": mse, "nll": nll, "combined": nll + w * mse}[kind]
    ghT, ghp = torch.autograd.grad(total, [hT, hp])
    return float(mse), float(nll), float(total), float(mae), ghT, ghp


@pytest.mark.parametrize("kind", ["mse", "nll", "combined"])
@pytest.mark.parametrize("shape", [(1, 100, 30, 1), (4, 100, 30, 4), (6, 25, 30, 2), (3, 7, 20, 1)])
def test_heads_loss_and_grads_vs_dense_reference(kind, shape):
    B_, S, n, wpg = shape
    H = 64
    hT, hp, y_stock, y_market, factor, inv_psi = _heads_case(B_, S, n, H, seed=B_ * 7 + S)
    mse, nll, total, mae, ghT, ghp = _heads_oracle(kind, 100.0, hT, hp, y_stock, y_market, factor, inv_psi, B_, S, wpg)
    lib = load()
    d = lambda t: t.to(DEV).contiguous()
    dh, dhp, dys, dym, dfa, dip = map(d, (hT, hp, y_stock, y_market, factor, inv_psi))
    losses = torch.zeros(4, device=DEV)
    dhT = torch.zeros_like(dh)
    hg = torch.zeros(2 * H + 2, device=DEV)
    al, be = torch.zeros(B_ * S, device=DEV), torch.zeros(B_ * S, device=DEV)
    ws = torch.empty(lib.b200_heads_workspace_bytes(B_, S, n, H), dtype=torch.uint8, device=DEV)
    check(lib.b200_heads_loss_fwd_bwd(B.LOSS_KINDS[kind], 100.0, ptr(dhp), H, ptr(dh), ptr(dys), ptr(dym), ptr(dfa), ptr(dip),
                                      B_, S, n, wpg, 1.0, ptr(al), ptr(be), ptr(losses), ptr(dhT), ptr(hg),
                                      ptr(ws), ws.numel(), None), "heads")
    got = losses.cpu().tolist()
    assert got[0] == pytest.approx(mse, rel=2e-5)
    assert got[3] == pytest.approx(mae, rel=2e-5)
    if kind != "mse":
        assert got[1] == pytest.approx(nll, rel=2e-5)
    assert got[2] == pytest.approx(total, rel=2e-5)
    assert rel_err(dhT.cpu().numpy(), ghT.numpy()) < 2e-4
    assert rel_err(hg.cpu().numpy(), ghp.numpy()) < 2e-4


def test_heads_rejects_bad_arguments():
    lib = load()
    t = torch.zeros(64, device=DEV)
    rc = lib.b200_heads_loss_fwd_bwd(0, 1.0, ptr(t), 64, ptr(t), ptr(t), ptr(t), None, None, 3, 100, 30, 2, 1.0,
                                     None, None, ptr(t), None, None, ptr(t), 64, None)
    assert rc == -1 and b"divisible" in lib.b200_last_error()
    rc = lib.b200_heads_loss_fwd_bwd(1, 1.0, ptr(t), 64, ptr(t), ptr(t), None, None, None, 2, 100, 30, 2, 1.0,
                                     None, None, ptr(t), None, None, ptr(t), 64, None)
    assert rc == -1 and b"NLL needs" in lib.b200_last_error()


@pytest.mark.parametrize("count,clip", [(51074, 5.0), (51076, 0.0), (1000003, 0.5), (7, 1.0)])
def test_clip_and_adam_vs_torch(count, clip):
    lib = load()
    g = torch.Generator().manual_seed(count)
    p0 = torch.randn(count, generator=g) * 0.1
    pad = (count + 3) // 4 * 4
    tp = torch.nn.Parameter(p0.clone())
    opt = torch.optim.Adam([tp], lr=1e-3, weight_decay=1e-5)
    dp, dg, dm, dv = (torch.zeros(pad, device=DEV) for _ in range(4))
    dp[:count] = p0.to(DEV)
    sq = torch.zeros(1, device=DEV)
    ws = torch.empty(lib.b200_sqnorm_workspace_bytes(pad), dtype=torch.uint8, device=DEV)
    for step in range(1, 4):
        grad = torch.randn(count, generator=g) * (3.0 if step == 2 else 0.01)
        tp.grad = grad.clone()
        if clip > 0:
            torch.nn.utils.clip_grad_norm_([tp], clip)
        opt.step()
        dg[:count] = grad.to(DEV)
        if clip > 0:
            check(lib.b200_grad_sqnorm(ptr(dg), pad, ptr(sq), ptr(ws), ws.numel(), None), "sqnorm")
            assert float(sq) == pytest.approx(float((grad.double() ** 2).sum()), rel=1e-5)
        check(lib.b200_adam_step(ptr(dp), ptr(dg), ptr(dm), ptr(dv), pad, 1e-3, 0.9, 0.999, 1e-8, 1e-5, step, clip,
                                 ptr(sq) if clip > 0 else None, 1, None), "adam")
        assert float(dg.abs().max()) == 0.0          # zero_grad
        np.testing.assert_allclose(dp[:count].cpu().numpy(), tp.detach().numpy(), rtol=2e-5, atol=2e-7)
    assert float(dp[count:].abs().sum()) == 0.0


@pytest.mark.parametrize("count,clip", [(51074, 5.0), (51076, 0.0), (117634, 0.5), (7, 1.0)])
def test_fused_clip_adam_vs_torch(count, clip):
    """b200_clip_adam_step (norm + clip + Adam in one launch, the path LstmEngine takes for the small and medium models) against
    clip_grad_norm_ + torch.optim.Adam for three steps; larger vectors are refused with UNSUPPORTED."""
    lib = load()
    g = torch.Generator().manual_seed(count + 1)
    p0 = torch.randn(count, generator=g) * 0.1
    pad = (count + 3) // 4 * 4
    tp = torch.nn.Parameter(p0.clone())
    opt = torch.optim.Adam([tp], lr=1e-3, weight_decay=1e-5)
    dp, dg, dm, dv = (torch.zeros(pad, device=DEV) for _ in range(4))
    dp[:count] = p0.to(DEV)
    sq = torch.zeros(1, device=DEV)
    for step in range(1, 4):
        grad = torch.randn(count, generator=g) * (3.0 if step == 2 else 0.01)
        tp.grad = grad.clone()
        if clip > 0:
            torch.nn.utils.clip_grad_norm_([tp], clip)
        opt.step()
        dg[:count] = grad.to(DEV)
        check(lib.b200_clip_adam_step(ptr(dp), ptr(dg), ptr(dm), ptr(dv), pad, 1e-3, 0.9, 0.999, 1e-8, 1e-5, step, clip, ptr(sq), None), "clip_adam")
        assert float(sq) == pytest.approx(float((grad.double() ** 2).sum()), rel=1e-5)
        assert torch.equal(dg[:count].cpu(), grad)                    # gradients are left untouched
        np.testing.assert_allclose(dp[:count].cpu().numpy(), tp.detach().numpy(), rtol=2e-5, atol=2e-7)
    assert float(dp[count:].abs().sum()) == 0.0
    big = torch.zeros(131072 + 4, device=DEV)
    assert lib.b200_clip_adam_step(ptr(big), ptr(big), ptr(big), ptr(big), big.numel(), 1e-3, 0.9, 0.999, 1e-8, 0.0, 1, 1.0, None, None) == -2


@pytest.mark.parametrize("variant,K,N", [(0, 16, 256), (0, 64, 256), (0, 128, 128), (1, 64, 128), (1, 256, 64),
                                         (1, 128, 128), (2, 128, 64), (2, 128, 256), (2, 64, 128)])
def test_tcgen05_gemm_selftest(variant, K, N):
    """UMMA descriptors (K-major and MN-major reads of the interleaved image), instruction descriptor and the
    TMEM load mapping: a 128xNxK product with the bf16x3 split must agree with the fp64 product to ~1e-5."""
    lib = load()
    g = torch.Generator().manual_seed(variant * 1000 + K + N)
    if variant == 0:
        A, Bm = torch.randn(128, K, generator=g), torch.randn(N, K, generator=g) * 0.125
        want = A.double() @ Bm.double().T
    elif variant == 1:
        A, Bm = torch.randn(128, K, generator=g), torch.randn(K, N, generator=g) * 0.125
        want = A.double() @ Bm.double()
    else:
        A, Bm = torch.randn(K, 128, generator=g), torch.randn(K, N, generator=g) * 0.125
        want = A.double().T @ Bm.double()
    D = torch.zeros(128, N, device=DEV)
    dA, dB = A.to(DEV), Bm.to(DEV)          # keep the device tensors alive across the call
    check(lib.b200_tc_selftest(ptr(dA), ptr(dB), variant, K, N, ptr(D), None), "tc_selftest")
    torch.cuda.synchronize()
    assert rel_err(D.cpu().numpy(), want.numpy()) < 2e-5
