"""-m gpu: the wide engine (hidden != 64; BASELINE.json configs[2]) -- its tcgen05 GEMM against fp64 matmuls, and the
LSTM training step (forward, objective, BPTT, clip, Adam) against the oracle, through the C-ABI."""
import numpy as np
import pytest
import torch

import lstm_oracle as O
import masters_thesis_b200 as B
from masters_thesis_b200._lib import check, load, ptr
from _gpu_util import DEV, golden_batches, make_engine, oracle_model, ref_batch, rel_err

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def data(golden_dir):
    return golden_batches(golden_dir)


@pytest.mark.parametrize("nsplit", [2, 1])
@pytest.mark.parametrize("variant,M,N,K", [(0, 300, 320, 192), (0, 128, 64, 64), (1, 300, 320, 192), (1, 77, 256, 1024),
                                           (2, 256, 128, 1000), (2, 1024, 256, 6400), (2, 128, 64, 33)])
def test_wide_gemm(nsplit, variant, M, N, K):
    """Ragged M (variants 0, 1), two N tiles (256 + 64), ragged and split K (variant 2), accumulate."""
    lib = load()
    g = torch.Generator().manual_seed(variant * 100 + M)
    a_shape = (K, M) if variant == 2 else (M, K)
    b_shape = (N, K) if variant == 0 else (K, N)
    A, Bm = torch.randn(a_shape, generator=g), torch.randn(b_shape, generator=g)
    ref = (A.double().T if variant == 2 else A.double()) @ (Bm.double().T if variant == 0 else Bm.double())
    Ad, Bd = A.to(DEV), Bm.to(DEV)
    C0 = torch.randn(M, N, generator=g)
    ws = torch.empty(64 * 1024 * 1024 // 4, dtype=torch.float32, device=DEV)
    tol = 2e-5 if nsplit == 2 else 1e-2       # relative to max|C|: three bf16 products carry 2^-17 per operand, one carries 2^-9
    scale = float(ref.abs().max())
    for acc in ((0, 1) if variant != 2 else (0,)):
        C = C0.to(DEV).clone()
        check(lib.b200_wide_gemm(nsplit, variant, ptr(Ad), ptr(Bd), ptr(C), M, N, K, acc, ptr(ws), ws.numel() * 4, None), "wide_gemm")
        want = ref + (C0.double() if acc else 0)
        assert float((C.cpu().double() - want).abs().max()) / scale < tol, (variant, acc)
    if variant != 2:      # with a workspace B went through pre-packed images + TMA bulk copies; without, the threads stage it
        C3 = torch.empty(M, N, device=DEV)
        check(lib.b200_wide_gemm(nsplit, variant, ptr(Ad), ptr(Bd), ptr(C3), M, N, K, 0, None, 0, None), "wide_gemm")
        assert float((C3.cpu().double() - ref).abs().max()) / scale < tol
    if variant == 2:      # without a workspace the same product runs unsplit; split-K is deterministic
        C1, C2 = torch.empty(M, N, device=DEV), torch.empty(M, N, device=DEV)
        check(lib.b200_wide_gemm(nsplit, 2, ptr(Ad), ptr(Bd), ptr(C1), M, N, K, 0, None, 0, None), "wide_gemm")
        check(lib.b200_wide_gemm(nsplit, 2, ptr(Ad), ptr(Bd), ptr(C2), M, N, K, 0, ptr(ws), ws.numel() * 4, None), "wide_gemm")
        # one CTA per tile accumulates all of K in TMEM: the tensor core's fp32 accumulation truncates, so a 6400-long K
        # drifts further than its split version (whose partials are summed with round-to-nearest)
        assert float((C1.cpu().double() - ref).abs().max()) / scale < 5 * tol
        assert torch.equal(C2, C)


def test_wide_gemm_rejects_bad_shapes():
    lib = load()
    t = torch.zeros(256 * 256, device=DEV)
    assert lib.b200_wide_gemm(2, 0, ptr(t), ptr(t), ptr(t), 128, 100, 64, 0, None, 0, None) != 0      # N % 64
    assert lib.b200_wide_gemm(2, 0, ptr(t), ptr(t), ptr(t), 128, 64, 100, 0, None, 0, None) != 0      # K % 64
    assert lib.b200_wide_gemm(2, 2, ptr(t), ptr(t), ptr(t), 100, 64, 64, 0, None, 0, None) != 0       # M % 128 (variant 2)
    assert lib.b200_wide_gemm(3, 0, ptr(t), ptr(t), ptr(t), 128, 64, 64, 0, None, 0, None) != 0
    assert b"wide_gemm" in lib.b200_last_error()


@pytest.mark.parametrize("H,layers,kind", [(128, 3, "combined"), (256, 2, "mse"), (256, 2, "nll")])
def test_wide_training_steps_vs_oracle(H, layers, kind, data):
    """Three optimisation steps of batch 2 (dropout 0): per-step loss <= 1e-4, step-0 gradients <= 1e-3 (whole vector and
    per tensor), parameters after the steps."""
    torch.manual_seed(21)
    m = O.OracleLstm(3, H, layers, 0.0)
    init = O.flat_params(m).clone()
    e = make_engine(init, layers, 0.0, "tc_precise", 3, H)
    opt = O.make_optimizer(m)
    order = [(0, 2), (2, 4), (0, 2)]
    for s, (lo, hi) in enumerate(order):
        batch = B.DeviceBatch.from_reference(ref_batch(data, lo, hi), DEV)
        e.forward(batch, training=True, dropout_seed=s)
        losses = e.heads(batch, kind, 100.0, None, 1.0, True).clone()
        e.backward(batch, True, dropout_seed=s)
        if s == 0:
            got = e.grads[: e.P].cpu().numpy().copy()
        e.optimizer_step(1e-4, 5.0)
        if s == 0:
            ls = O.batch_losses(m, ref_batch(data, lo, hi), kind, 100.0)
            opt.zero_grad()
            ls["total"].backward()
            want = O.flat_grads(m).numpy().copy()
            assert rel_err(got, want) < 1e-3
            off = 0
            for name, p_ in m.named_parameters():
                n = p_.numel()
                assert rel_err(got[off:off + n], want[off:off + n]) < 1e-3, name
                off += n
            O.clip_global_norm(list(m.parameters()), 5.0)
            opt.step()
            total = float(ls["total"].detach())
        else:
            total = O.train_step(m, opt, ref_batch(data, lo, hi), kind, 5.0)[0]["total"]
        assert float(losses[2]) == pytest.approx(total, rel=1e-4), s
    diff = np.abs(e.params[: e.P].cpu().numpy() - O.flat_params(m).numpy())
    assert np.mean(diff > 2e-6) < 1e-3 and diff.max() < 1e-4


def test_wide_dropout_step_vs_oracle_with_shared_masks(data):
    """The widened model's dropout (0.3) with the kernels' Philox masks injected into the oracle."""
    H, layers, p, seed = 256, 3, 0.3, 0x5EED1234
    torch.manual_seed(12)
    m = O.OracleLstm(3, H, layers, 0.0)
    e = make_engine(O.flat_params(m), layers, p, "tc_precise", 3, H)
    batch = B.DeviceBatch.from_reference(ref_batch(data, 1, 3), DEV)
    e.forward(batch, training=True, dropout_seed=seed)
    losses = e.heads(batch, "combined", 100.0, None, 1.0, True).clone()
    e.backward(batch, True, dropout_seed=seed)
    masks = [torch.from_numpy(O.dropout_mask(seed, l + 1, 200, 60, H, p)).permute(1, 0, 2) for l in range(layers - 1)]
    em = O.ExplicitMaskLstm(m)
    ls = O.batch_losses(m, ref_batch(data, 1, 3), "combined", 100.0, forward=lambda x: em(x, masks))
    m.zero_grad()
    ls["total"].backward()
    assert float(losses[2]) == pytest.approx(float(ls["total"]), rel=1e-4)
    assert rel_err(e.grads[: e.P].cpu().numpy(), O.flat_grads(m).numpy()) < 1e-3


def test_wide_fast_mode_and_eval(data):
    """configs[2] as named: plain bf16 GEMMs with fp32 accumulation (tc_fast) -- loss within 1 %, gradient direction
    preserved; evaluation (batch-1 semantics) on the precise engine equals the oracle's."""
    H, layers = 256, 2
    torch.manual_seed(4)
    m = O.OracleLstm(3, H, layers, 0.0)
    flat = O.flat_params(m).clone()
    ref = ref_batch(data, 0, 2)
    ls = O.batch_losses(m, ref, "mse", 100.0)
    ls["total"].backward()
    want = O.flat_grads(m).numpy()
    e = make_engine(flat, layers, 0.0, "tc_fast", 3, H)
    batch = B.DeviceBatch.from_reference(ref, DEV)
    e.forward(batch, training=True)
    losses = e.heads(batch, "mse", 100.0, None, 1.0, True).clone()
    e.backward(batch, True)
    got = e.grads[: e.P].cpu().numpy()
    assert float(losses[2]) == pytest.approx(float(ls["total"]), rel=1e-2)
    cos = float(np.dot(got, want) / (np.linalg.norm(got) * np.linalg.norm(want)))
    assert cos > 0.999
    ep = make_engine(flat, layers, 0.0, "tc_precise", 3, H)
    v = ep.eval_windows(B.DeviceBatch.from_reference(ref_batch(data, 4, 6), DEV), "combined")
    ov = O.eval_windows(m, *ref_batch(data, 4, 6), "combined")
    assert float(v[2]) == pytest.approx(ov["total"], rel=1e-4)


def test_wide_config_through_trainer():
    """`model=wide` (hidden 256, 8 layers, dropout 0.3) runs through the trainer object that train.py builds."""
    torch.manual_seed(1234)
    model = B.FinancialLstmMse(3, 256, 8, 0.3, 1e-4)
    dm = B.FinancialLstmDataModule(lookback_window=60, target_window=30, stride=90, batch_size=4, seed=9, n_samples=90 * 30,
                                   device=DEV)
    tr = B.B200Trainer(max_epochs=1, gradient_clip_val=5.0, max_steps=3, engine="auto", precision="bf16-mixed", verbose=False)
    tr.record_steps = True
    tr.fit(model, dm)
    losses = [float(l[2]) for l in tr.step_losses]
    assert len(losses) == 3 and all(np.isfinite(losses))
    assert np.isfinite(float(tr.checkpoint_callback.best_model_score))


def test_wide_fused_cell_path_matches_unfused(data):
    """b200_set_option("wide_fuse", 1) (cell update in the epilogue of the recurrent GEMM, gate-interleaved weight images): same hidden
    states and gradients as the default path."""
    H, layers = 256, 2
    torch.manual_seed(8)
    flat = O.flat_params(O.OracleLstm(3, H, layers, 0.0)).clone()
    batch = B.DeviceBatch.from_reference(ref_batch(data, 0, 2), DEV)
    outs = []
    lib = B._lib.load()
    for fuse in (0, 1):
        assert lib.b200_set_option(b"wide_fuse", fuse) == 0
        e = make_engine(flat, layers, 0.0, "tc_precise", 3, H)
        e.forward(batch, training=True)
        losses = e.heads(batch, "combined", 100.0, None, 1.0, True).clone()
        e.backward(batch, True)
        outs.append((e.stash_h.clone(), losses, e.grads[: e.P].clone()))
    lib.b200_set_option(b"wide_fuse", 0)
    assert float((outs[0][0] - outs[1][0]).abs().max()) < 2e-6
    assert float(outs[0][1][2]) == pytest.approx(float(outs[1][1][2]), rel=1e-6)
    assert rel_err(outs[1][2].cpu().numpy(), outs[0][2].cpu().numpy()) < 1e-5


# ------------------------------------------------------------------------------------ bf16 operand-image engine (lstm_wide2.cu)
def _oracle_step(m, ref, kind, masks=None):
    em = O.ExplicitMaskLstm(m)
    ls = O.batch_losses(m, ref, kind, 100.0, forward=(lambda x: em(x, masks)) if masks is not None else None)
    m.zero_grad()
    ls["total"].backward()
    with torch.no_grad():
        seqs = em(ref[0].flatten(0, 1), masks, return_all=True)[2]
    return float(ls["total"].detach()), O.flat_grads(m).numpy().copy(), [s.permute(1, 0, 2).numpy() for s in seqs]


@pytest.mark.parametrize("H,layers,p,nwin,kind", [(256, 2, 0.0, 2, "mse"), (128, 3, 0.0, 3, "combined"), (256, 3, 0.3, 3, "mse"),
                                                  (256, 8, 0.3, 2, "mse")])
def test_wide_bf16_engine_vs_oracle(H, layers, p, nwin, kind, data):
    """B200_MODE_TC_FAST at hidden 128 / 256 (bf16 operands and bf16 activation stash, fp32 accumulation and cell state), including
    the configs[2] shape hidden 256 x 8 layers with dropout 0.3: every layer's hidden states within 1e-2 absolute of the fp32
    oracle (bf16 has 8 mantissa bits), loss within 1e-3 relative, every parameter tensor's gradient within 2e-2 of the tensor's
    largest gradient, direction of the whole gradient cosine > 0.9999.  Dropout through the shared Philox masks."""
    seed = 0xB16B00
    torch.manual_seed(31 + layers)
    m = O.OracleLstm(3, H, layers, 0.0)
    flat = O.flat_params(m).clone()
    ref = ref_batch(data, 0, nwin)
    masks = None
    if p > 0:
        masks = [torch.from_numpy(O.dropout_mask(seed, l + 1, nwin * 100, 60, H, p)).permute(1, 0, 2) for l in range(layers - 1)]
    total, want, seqs = _oracle_step(m, ref, kind, masks)
    e = make_engine(flat, layers, p, "tc_fast", 3, H)
    batch = B.DeviceBatch.from_reference(ref, DEV)
    e.forward(batch, training=True, dropout_seed=seed)
    hs = [e.hidden_states(l, nwin).cpu().numpy() for l in range(layers)]
    losses = e.heads(batch, kind, 100.0, None, 1.0, True).clone()
    e.backward(batch, True, dropout_seed=seed)
    got = e.grads[: e.P].cpu().numpy()
    for l in range(layers):
        np.testing.assert_allclose(hs[l], seqs[l], rtol=0, atol=1e-2, err_msg=f"layer {l}")
    assert float(losses[2]) == pytest.approx(total, rel=1e-3)
    cos = float(np.dot(got, want) / (np.linalg.norm(got) * np.linalg.norm(want)))
    assert cos > 0.9999
    off = 0
    for name, p_ in m.named_parameters():
        n = p_.numel()
        assert rel_err(got[off:off + n], want[off:off + n]) < 2e-2, name
        off += n


def test_wide_bf16_engine_is_deterministic_and_micro_batches_add_up(data):
    """Two runs give bit-identical gradients; a 4-window step equals the sum of its two 2-window halves to fp32 summation order
    (what ChunkedBatch + b200_grad_accumulate rely on)."""
    H, layers = 256, 2
    torch.manual_seed(77)
    flat = O.flat_params(O.OracleLstm(3, H, layers, 0.0)).clone()
    e = make_engine(flat, layers, 0.0, "tc_fast", 3, H)

    def run(lo, hi, scale):
        batch = B.DeviceBatch.from_reference(ref_batch(data, lo, hi), DEV)
        e.forward(batch, training=True)
        e.heads(batch, "mse", 100.0, None, scale, True)
        e.backward(batch, True)
        return e.grads[: e.P].clone()

    whole = run(0, 4, 1.0)
    again = run(0, 4, 1.0)
    assert torch.equal(whole, again)
    halves = run(0, 2, 0.5) + run(2, 4, 0.5)
    assert rel_err(halves.cpu().numpy(), whole.cpu().numpy()) < 2e-4


def test_wide_images_option_selects_the_fp32_storage_path(data):
    """b200_set_option("wide_images", 0) sends bf16 mode through lstm_wide.cu (fp32 activations in global memory): same loss to
    bf16 accuracy, so the two implementations check each other."""
    H, layers = 256, 2
    torch.manual_seed(3)
    flat = O.flat_params(O.OracleLstm(3, H, layers, 0.0)).clone()
    batch = B.DeviceBatch.from_reference(ref_batch(data, 0, 2), DEV)
    lib = B._lib.load()
    out = []
    try:
        for images in (1, 0):
            assert lib.b200_set_option(b"wide_images", images) == 0
            e = make_engine(flat, layers, 0.0, "tc_fast", 3, H)
            e.forward(batch, training=True)
            losses = e.heads(batch, "mse", 100.0, None, 1.0, True).clone()
            e.backward(batch, True)
            out.append((float(losses[2]), e.grads[: e.P].cpu().numpy().copy()))
    finally:
        lib.b200_set_option(b"wide_images", 1)
    assert out[0][0] == pytest.approx(out[1][0], rel=1e-3)
    assert rel_err(out[0][1], out[1][1]) < 2e-2


def test_wide_forward_cluster_multicast_option_is_bit_identical(data):
    """b200_set_option("wide_mc", 1): the forward kernel as two-CTA clusters that multicast the activation chunks to each other (all
    sources are this GPU's own workspace) computes exactly what the default kernel computes."""
    H, layers, p, seed = 256, 3, 0.3, 99
    torch.manual_seed(5)
    flat = O.flat_params(O.OracleLstm(3, H, layers, 0.0)).clone()
    batch = B.DeviceBatch.from_reference(ref_batch(data, 0, 3), DEV)
    lib = B._lib.load()
    out = []
    try:
        for mc in (0, 1):
            assert lib.b200_set_option(b"wide_mc", mc) == 0
            e = make_engine(flat, layers, p, "tc_fast", 3, H)
            e.forward(batch, training=True, dropout_seed=seed)
            hs = [e.hidden_states(l, 3).clone() for l in range(layers)]
            losses = e.heads(batch, "mse", 100.0, None, 1.0, True).clone()
            e.backward(batch, True, dropout_seed=seed)
            out.append((hs, losses, e.grads[: e.P].clone()))
    finally:
        lib.b200_set_option(b"wide_mc", 0)
    for a, b in zip(out[0][0], out[1][0]):
        assert torch.equal(a, b)
    assert torch.equal(out[0][1], out[1][1]) and torch.equal(out[0][2], out[1][2])


def test_wide_bf16_training_steps_track_the_oracle(data):
    """Six optimisation steps (clip 5.0, Adam 1e-4) of the hidden-256 x 3-layer model in bf16 mode against the fp32 oracle loop on the
    same batches: per-step loss within 2e-3 relative; parameters after the steps: six Adam steps move a parameter by at most 6e-4 and the
    sign of a near-zero gradient may differ in bf16, so the bound on a single parameter is 1.2e-3, the mean difference stays below 5e-6
    and fewer than 2 % of the parameters differ by more than 5e-5."""
    H, layers = 256, 3
    torch.manual_seed(41)
    m = O.OracleLstm(3, H, layers, 0.0)
    e = make_engine(O.flat_params(m).clone(), layers, 0.0, "tc_fast", 3, H)
    opt = O.make_optimizer(m)
    order = [(0, 2), (2, 4), (4, 6), (0, 2), (2, 4), (4, 6)]
    for s, (lo, hi) in enumerate(order):
        ref = ref_batch(data, lo, hi)
        batch = B.DeviceBatch.from_reference(ref, DEV)
        e.forward(batch, training=True, dropout_seed=s)
        losses = e.heads(batch, "combined", 100.0, None, 1.0, True).clone()
        e.backward(batch, True, dropout_seed=s)
        e.optimizer_step(1e-4, 5.0)
        total = O.train_step(m, opt, ref, "combined", 5.0)[0]["total"]
        assert float(losses[2]) == pytest.approx(total, rel=2e-3), s
    diff = np.abs(e.params[: e.P].cpu().numpy() - O.flat_params(m).numpy())
    assert diff.max() < 1.2e-3 and diff.mean() < 5e-6 and np.mean(diff > 5e-5) < 0.02, (diff.max(), diff.mean(), np.mean(diff > 5e-5))
