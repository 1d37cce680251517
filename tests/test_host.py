"""CPU tests (no GPU): library loads and exports every symbol of include/b200_lstm.h, config
composition, model mirror initialisation, datamodule split arithmetic, gradient bucketing under gloo."""
import os
import re
import subprocess
import sys
from pathlib import Path

import numpy as np
import pytest
import torch

import masters_thesis_b200 as B

ROOT = Path(__file__).resolve().parent.parent
CONFIGS = ROOT / "configs"


def test_library_exports_every_declared_symbol():
    hdr = (ROOT / "include" / "b200_lstm.h").read_text()
    declared = set(re.findall(r"B200_API [\w\s\*]+?\b(b200_\w+)\(", hdr))
    assert len(declared) >= 19
    lib = B._lib.load()
    for name in declared:
        assert hasattr(lib, name), name
    assert declared == set(B._lib.PROTOTYPES), declared ^ set(B._lib.PROTOTYPES)
    # host-only entry points work without a GPU
    assert lib.b200_version() >= 100
    assert lib.b200_param_count(3, 64, 2) == 51074
    assert lib.b200_param_count(3, 64, 4) == 117634
    assert lib.b200_param_count(3, 64, 8) == 250754
    assert lib.b200_param_offset(3, 64, 2, 2) == 51074 - 130


def test_missing_library_fails_loudly(tmp_path):
    with pytest.raises(B.B200Error, match="no CPU fallback"):
        B._lib.load(tmp_path / "nope.so")


def test_engine_refuses_to_run_without_cuda():
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(B.B200Error):
        B.LstmEngine(3, 64, 2)


def test_config_defaults_and_overrides():
    cfg = B.compose_config(CONFIGS)
    assert cfg.datamodule.name == "synthetic" and cfg.model.name == "small" and cfg.loss.name == "mse"
    assert cfg.trainer.name == "fast" and cfg.trainer.max_epochs == 4 and cfg.trainer.gradient_clip_val == 5.0
    assert cfg.model.input_size == 3 and cfg.model.hidden_size == 64 and cfg.model.num_layers == 2
    assert cfg.logger.name == "FinancialLstm/synthetic"
    assert cfg.logger.version == "mse_small_lr0.0001_fast"
    assert cfg.checkpoint is None
    cfg = B.compose_config(CONFIGS, overrides=["loss=combined", "model=large", "trainer=b200", "loss.mse_weight=0.7",
                                                "datamodule.interaction_only=false", "datamodule.batch_size=4096"])
    assert cfg.loss.module_class == "FinancialLstmCombined" and cfg.loss.mse_weight == 0.7
    assert cfg.model.num_layers == 8 and cfg.model.input_size == 5
    assert cfg.trainer.name == "b200" and cfg.datamodule.batch_size == 4096
    assert cfg.logger.version == "combined_large_lr0.0001_b200"
    with pytest.raises(ValueError, match="Could not find 'trainer/slower'"):
        B.compose_config(CONFIGS, overrides=["trainer=slower"])


def test_multirun_grid():
    from masters_thesis_b200.config import expand_multirun, split_cli
    multi, ov = split_cli(["-m", "datamodule=synthetic", "model=small,medium,large", "loss=mse,nll,combined", "trainer=fast,slow"])
    assert multi
    grid = expand_multirun(ov)
    assert len(grid) == 18 and grid[0] == ["datamodule=synthetic", "model=small", "loss=mse", "trainer=fast"]


def test_model_mirror_initialises_like_torch_modules():
    torch.manual_seed(1234)
    m = B.FinancialLstmNll(3, 64, 2, 0.2, 1e-4)
    torch.manual_seed(1234)
    lstm = torch.nn.LSTM(3, 64, 2, dropout=0.2, batch_first=True)
    a, b = torch.nn.Linear(64, 1), torch.nn.Linear(64, 1)
    want = torch.cat([p.detach().reshape(-1) for p in list(lstm.parameters()) + list(a.parameters()) + list(b.parameters())])
    assert torch.equal(m.flat_parameters(), want)
    sd = m.state_dict()
    assert list(sd)[:4] == ["lstm.weight_ih_l0", "lstm.weight_hh_l0", "lstm.bias_ih_l0", "lstm.bias_hh_l0"]
    assert sd["alpha_head.weight"].shape == (1, 64) and sd["beta_head.bias"].shape == (1,)
    m2 = B.FinancialLstmNll(3, 64, 2, 0.2, 1e-4)
    m2.load_state_dict(sd)
    assert torch.equal(m2.flat_parameters(), want)
    assert B.get_model_class("FinancialLstmCombined")(mse_weight=3.0).hparams["mse_weight"] == 3.0
    with pytest.raises(ValueError):
        B.get_model_class("Nope")


def test_initial_state_matches_reference_golden(golden_dir):
    g = np.load(golden_dir / "golden_mse.npz")
    torch.manual_seed(int(g["meta"][1]))
    m = B.FinancialLstmMse(3, 64, 2, 0.0, 1e-4)
    assert np.array_equal(m.flat_parameters().numpy(), g["init"])
    assert m.names == list(g["names"])


def test_datamodule_split_matches_reference_arithmetic():
    dm = B.FinancialLstmDataModule(lookback_window=60, target_window=30, stride=90, batch_size=1)
    dm.prepare_data()
    assert (dm.n_windows, dm.train_end, dm.val_end) == (11111, 7777, 9999)
    assert dm.n_windows - dm.val_end == 1112
    with pytest.raises(ValueError):
        B.FinancialLstmDataModule(lookback_window=10, target_window=20, prediction_task=False)


class _FakeSource:
    device = "cpu"

    def generate(self, window_idx=None, first_window=0, n_windows=None, checked=False, micro=None, out=None):
        return window_idx.tolist() if window_idx is not None else list(range(first_window, first_window + n_windows))


def test_loader_rank_shares_follow_distributed_sampler():
    """Under data parallelism batch_size is per rank (Lightning DDP on the reference, train.py:169) and every rank sees its
    DistributedSampler share of the epoch: the reference default batch_size 1 with WORLD_SIZE 2 trains on every window
    (nothing skipped, nothing dropped), ragged epochs are padded by wrapping around."""
    from masters_thesis_b200.data import _Loader, _StreamLoader, rank_share
    n, world = 11, 2
    seen = []
    for rank in range(world):
        gen = torch.Generator().manual_seed(3)
        batches = list(_Loader(_FakeSource(), 100, 100 + n, 1, True, gen, rank, world))
        assert len(batches) == 6 == len(_Loader(_FakeSource(), 100, 100 + n, 1, True, gen, rank, world))
        assert all(len(b) == 1 for b in batches)
        seen.append([b[0] for b in batches])
    order = (torch.randperm(n, generator=torch.Generator().manual_seed(3)) + 100).tolist()
    padded = order + order[:1]
    assert seen[0] == padded[0::2] and seen[1] == padded[1::2]
    assert set(seen[0]) | set(seen[1]) == set(range(100, 100 + n))
    # torch's own sampler agrees (shuffle off: same order in, same share out)
    from torch.utils.data.distributed import DistributedSampler
    for rank in range(3):
        want = list(DistributedSampler(range(10), num_replicas=3, rank=rank, shuffle=False))
        assert rank_share(torch.arange(10), rank, 3).tolist() == want
    # batch_size 4 per rank: ceil(6 / 4) = 2 batches, the last one ragged
    b = list(_Loader(_FakeSource(), 0, n, 4, False, None, 1, 2))
    assert [len(x) for x in b] == [4, 2] and b[0] == [1, 3, 5, 7]
    # streaming: step k of rank r = windows first + (k * W + r) * B .. + B, disjoint over ranks and steps
    s0 = list(_StreamLoader(_FakeSource(), 1000, 8, 2, 0, 2))
    s1 = list(_StreamLoader(_FakeSource(), 1000, 8, 2, 1, 2))
    assert s0[0] == list(range(1000, 1008)) and s1[0] == list(range(1008, 1016)) and s0[1] == list(range(1016, 1024))
    with pytest.raises(ValueError):
        _Loader(_FakeSource(), 0, 4, 0, False)


def test_stale_library_is_detected(tmp_path, monkeypatch):
    """_lib.load() compares the hash compiled into the binary with the sources next to it."""
    from masters_thesis_b200 import build as builder
    lib = B._lib.load()
    assert lib.b200_build_hash().decode() == builder.source_hash(builder._extra_flags())
    monkeypatch.setattr(builder, "source_hash", lambda extra=(): "0" * 16)
    with pytest.raises(B.B200Error, match="built from other sources"):
        B._lib.load(B._lib.LIB_PATH)


def test_train_entry_names_b200_for_reference_trainers(capsys):
    """`python train.py` with the reference's untouched defaults (trainer: fast) is the reference's Lightning path: without
    lightning/hydra and a reference checkout it must say so and name trainer=b200, not train something else."""
    sys.path.insert(0, str(ROOT))
    import importlib
    import importlib.util
    if importlib.util.find_spec("lightning") is not None:
        pytest.skip("lightning present: the fallthrough would run the reference")
    train = importlib.import_module("train")
    with pytest.raises(SystemExit) as ei:
        train.main([])
    assert "trainer=b200" in str(ei.value) and "trainer=fast" in str(ei.value)
    for name, (epochs, clip, cadence) in {"b200_fast": (4, 5.0, None), "b200_slow": (32, 2.0, 4), "b200_slowest": (1024, 1.0, 64)}.items():
        cfg = B.compose_config(CONFIGS, overrides=[f"trainer={name}"])
        ref = B.compose_config(CONFIGS, overrides=[f"trainer={name[5:]}"])
        assert train.is_b200_trainer(cfg) and not train.is_b200_trainer(ref)
        assert (cfg.trainer.max_epochs, cfg.trainer.gradient_clip_val, cfg.trainer.get("check_val_every_n_epoch")) == (epochs, clip, cadence)
        assert (ref.trainer.max_epochs, ref.trainer.gradient_clip_val, ref.trainer.get("check_val_every_n_epoch")) == (epochs, clip, cadence)


def test_plateau_schedule_matches_torch():
    from masters_thesis_b200.trainer import PlateauSchedule
    p = torch.nn.Parameter(torch.zeros(1))
    opt = torch.optim.Adam([p], lr=1e-4)
    ref = torch.optim.lr_scheduler.ReduceLROnPlateau(opt, mode="min", factor=0.5, patience=2)
    mine = PlateauSchedule(1e-4)
    for v in [5.0, 4.0, 4.1, 4.2, 4.3, 4.0, 3.99999, 4.5, 4.6, 4.7, 4.8, 4.9, 1.0, 2.0, 2.0, 2.0, 2.0]:
        ref.step(v)
        assert mine.step(v) == pytest.approx(opt.param_groups[0]["lr"], rel=1e-12)


_WORKER = r'''
import os, sys, torch, torch.distributed as dist
sys.path.insert(0, sys.argv[1])
from masters_thesis_b200.grad_sync import GradBucketer
dist.init_process_group("gloo", init_method="tcp://127.0.0.1:" + sys.argv[2], rank=int(sys.argv[3]), world_size=2)
rank = dist.get_rank()
P, slices = 1000, [(900, 100), (600, 300), (250, 350), (0, 250)]       # heads, then layers top-down
for bucket_mb, expect in [(0.0, 4), (400 * 4 / 2**20, 2), (100.0, 1)]:
    g = torch.arange(P, dtype=torch.float32) * (rank + 1) / 2           # pre-scaled by 1/W
    s = GradBucketer(2, bucket_mb)
    s.begin(g, P)
    for i, (o, n) in enumerate(slices):
        s.layer_ready(o, n, last=(i == len(slices) - 1))
    s.wait()
    want = torch.arange(P, dtype=torch.float32) * 1.5                   # mean of rank grads x and 2x
    assert torch.allclose(g, want), (bucket_mb, (g - want).abs().max())
    assert len(s.flushes) == expect, (bucket_mb, s.flushes)
    assert sum(n for _, n in s.flushes) == P
try:
    s.begin(g, P); s.layer_ready(900, 100); s.layer_ready(0, 250)
    raise SystemExit("non-contiguous slice accepted")
except ValueError:
    pass
dist.barrier()
print("ok", rank)
'''


def test_grad_bucketer_world2_gloo(tmp_path):
    script = tmp_path / "w.py"
    script.write_text(_WORKER)
    port = str(29500 + os.getpid() % 2000)
    procs = [subprocess.Popen([sys.executable, str(script), str(ROOT), port, str(r)], stdout=subprocess.PIPE,
                              stderr=subprocess.STDOUT, text=True) for r in range(2)]
    outs = [p.communicate(timeout=120)[0] for p in procs]
    for p, o in zip(procs, outs):
        assert p.returncode == 0, o
        assert "ok" in o


def test_sweep_plan_assigns_one_gpu_per_job():
    sys.path.insert(0, str(ROOT / "sweeps"))
    import run_sweep
    grid = ["datamodule=synthetic", "model=small,medium,large", "loss=mse,nll,combined", "trainer=b200_fast,b200_slow"]
    assert len(run_sweep.plan(grid, 8)) == 18           # the reference's sweeps/experiment_synthetic.sh grid
    jobs = run_sweep.plan(["datamodule=synthetic", "model=small,medium,large", "loss=mse,nll,combined", "trainer=b200"], 8)
    assert len(jobs) == 9
    assert [j["gpu"] for j in jobs] == [0, 1, 2, 3, 4, 5, 6, 7, 0]
    assert jobs[4]["overrides"] == ["datamodule=synthetic", "model=medium", "loss=nll", "trainer=b200"]


def test_fama_french_loader(tmp_path):
    """load_fama_french_25 (mirror of FamaFrench25Portfolios.load, src/data.py:86-123) on CSVs of the published layout:
    header lines, 3125 skipped old rows, percent returns, -99.99 / -999 missing markers."""
    rng = np.random.default_rng(0)
    n_old, n_new = 3125, 40
    dates = np.arange(19260701, 19260701 + n_old + n_new + 1)
    mkt, rf = rng.normal(0.03, 1.0, dates.size).round(2), np.full(dates.size, 0.01)
    p25 = rng.normal(0.04, 1.2, (dates.size, 25)).round(2)
    p25[n_old + 1 + 7, 3] = -99.99                    # one missing day in the kept part ...
    rf[n_old + 1 + 7] = 0.0                           # ... on a zero-RF day: the reference tests the marker AFTER subtracting RF (src/data.py:110-113)
    with open(tmp_path / "F-F_Research_Data_Factors_daily.csv", "w") as f:
        f.write("line1\nline2\nline3\n\n")             # ff3_skip = 4 lines before the column header
        for i, dt in enumerate(dates):
            f.write(f"{dt},{mkt[i]:.2f},0.00,0.00,{rf[i]:.2f}\n")
    with open(tmp_path / "25_Portfolios_5x5_Daily.csv", "w") as f:
        f.write("\n" * 18)
        for i, dt in enumerate(dates):
            f.write(f"{dt}," + ",".join(f"{v:.2f}" for v in p25[i]) + "\n")
    stocks, market = B.data.load_fama_french_25(tmp_path)
    keep = np.ones(n_new, bool); keep[7] = False
    m = mkt[n_old + 1:][keep].astype(np.float32)
    p = (p25[n_old + 1:] - rf[n_old + 1:, None]).astype(np.float32)[keep]
    assert stocks.shape == (25, n_new - 1) and market.shape == (n_new - 1,)
    np.testing.assert_allclose(market.numpy(), 100 * (np.log(m + 100) - np.log(np.float32(100))), rtol=1e-4, atol=3e-4)   # fp32 cancellation in log(x+100)-log(100)
    np.testing.assert_allclose(stocks.numpy(), (100 * (np.log(p + 100) - np.log(np.float32(100)))).T, rtol=1e-4, atol=3e-4)
    ref_root = Path("/root/reference")
    if ref_root.exists():                             # build container only: the reference's own loader on the same files
        import ref_stubs
        src = ref_stubs.install()
        p_ref, m_ref = src.data.FamaFrench25Portfolios.load(tmp_path)
        np.testing.assert_allclose(stocks.numpy(), p_ref.numpy(), rtol=1e-6, atol=1e-6)
        np.testing.assert_allclose(market.numpy(), m_ref.numpy(), rtol=1e-6, atol=1e-6)


def test_wide_workspace_sizes_are_pure_arithmetic():
    """b200_lstm_workspace_bytes needs no GPU: the bf16 wide engine (operand images, lstm_wide2.cu) needs less workspace per row than the
    fp32-storage path (b200_set_option("wide_images", 0)), both grow linearly with the rows, and unknown options are rejected."""
    from masters_thesis_b200 import _lib
    lib = _lib.load()
    fast, precise = _lib.MODES["tc_fast"], _lib.MODES["tc_precise"]
    T, F, H, L = 60, 3, 256, 8
    img = [int(lib.b200_lstm_workspace_bytes(fast, n, T, F, H, L)) for n in (12800, 25600, 51200)]
    try:
        assert lib.b200_set_option(b"wide_images", 0) == 0
        f32 = [int(lib.b200_lstm_workspace_bytes(fast, n, T, F, H, L)) for n in (12800, 25600, 51200)]
    finally:
        lib.b200_set_option(b"wide_images", 1)
    assert all(a > 0 for a in img) and all(a < b for a, b in zip(img, f32))
    # linear in the rows up to the constant part (packed weights, split-K partials)
    assert abs((img[2] - img[1]) - 2 * (img[1] - img[0])) < 0.01 * img[2]
    per_row_step_layer = (img[2] - img[1]) / (25600 * T * L)
    assert 3000 < per_row_step_layer < 5000            # gates 2 KB + h and masked h 1 KB + c 1 KB + dx
    assert int(lib.b200_lstm_workspace_bytes(precise, 51200, T, F, H, L)) == f32[2]
    assert lib.b200_set_option(b"no_such_option", 1) != 0
