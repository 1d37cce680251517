"""Device-resident mirror of the reference datamodule (src/data.py:133-250).

Same constructor arguments and the same prepare_data()/setup()/..._dataloader() surface, but no
dataset file is materialised: every window is a pure function of (seed, window index) evaluated by
the `b200_gen_batch` kernel (generator spec: oracle/lstm_oracle.py), so a batch is produced on the
device from a list of window indices.  The reference's 70/20/10 contiguous split of the
(n_samples - 90)//stride + 1 windows (src/data.py:221-234) and its shuffled train loader
(src/data.py:236-238) are kept; `streaming=True` drops the epoch structure and walks fresh windows
(needed once batch_size exceeds the 7,777 training windows of the default 1,000,000-day series).

When `data_dir` holds the reference's `stocks.pt` + `market.pt` (src/data.py:193-196; `datamodule=real`, or a
synthetic series somebody saved), the series is uploaded once and `b200_series_batch` cuts the same windows out
of it on the device.  `prediction_task=false` (reconstruction split, src/common.py:107-110) runs through the same
kernel; with the generator its series is first materialised on the device by `b200_gen_series`.
"""
from __future__ import annotations

from dataclasses import dataclass
from pathlib import Path

import torch

from . import _lib
from ._lib import check, ptr


@dataclass
class DeviceBatch:
    """One collated batch in the kernels' layout (include/b200_lstm.h)."""
    x: torch.Tensor          # [T][N][F]
    y_stock: torch.Tensor    # [N][n]
    y_market: torch.Tensor   # [B][n]
    factor: torch.Tensor     # [B][2]
    inv_psi: torch.Tensor    # [N]
    n_windows: int
    n_stocks: int
    ols_alpha: torch.Tensor | None = None     # [N] per-window OLS coefficients of the target (series sources, on request)
    ols_beta: torch.Tensor | None = None

    def to_reference(self, alphas=None, betas=None):
        """The reference's collated batch [X (B,S,T,F), y (B,S,n,4), factor (B,2), inv_psi (B,S)]
        (src/data.py:216, 236-238); channels 2,3 of y carry the true alpha/beta when given."""
        B, S = self.n_windows, self.n_stocks
        T, N, F = self.x.shape
        n = self.y_stock.shape[1]
        X = self.x.permute(1, 0, 2).reshape(B, S, T, F)
        ys = self.y_stock.reshape(B, S, n)
        ym = self.y_market[:, None, :].expand(B, S, n)
        a = torch.zeros(S, device=self.x.device) if alphas is None else alphas.to(self.x.device)
        b = torch.zeros(S, device=self.x.device) if betas is None else betas.to(self.x.device)
        y = torch.stack([ys, ym, a[None, :, None].expand(B, S, n), b[None, :, None].expand(B, S, n)], dim=-1)
        return [X.contiguous(), y.contiguous(), self.factor.clone(), self.inv_psi.reshape(B, S).clone()]

    @staticmethod
    def from_reference(batch, device):
        """Inverse of to_reference(): accepts the list of four tensors the reference's loaders yield.  On a CUDA device
        the layout change is one launch of the library's own kernel (b200_collate_reference)."""
        X, y, factor, inv_psi = [t.to(device=device, dtype=torch.float32) for t in batch]
        B, S, T, F = X.shape
        n = y.shape[2]
        if X.is_cuda and y.shape[3] == 4:
            X, y = X.contiguous(), y.contiguous()
            e = lambda *s: torch.empty(*s, dtype=torch.float32, device=X.device)
            out = DeviceBatch(e(T, B * S, F), e(B * S, n), e(B, n), factor.contiguous(), inv_psi.reshape(B * S).contiguous(), B, S)
            stream = torch.cuda.current_stream(X.device).cuda_stream
            check(_lib.load().b200_collate_reference(ptr(X), ptr(y), B, S, T, n, F, ptr(out.x), ptr(out.y_stock), ptr(out.y_market),
                                                     stream), "b200_collate_reference")
            return out
        x = X.reshape(B * S, T, F).permute(1, 0, 2).contiguous()
        return DeviceBatch(x, y[..., 0].reshape(B * S, n).contiguous(), y[:, 0, :, 1].contiguous(),
                           factor.contiguous(), inv_psi.reshape(B * S).contiguous(), B, S)


class DeviceBatchSource:
    """Allocates (and reuses) the output tensors of b200_gen_batch."""

    def __init__(self, device, seed=1234, n_stocks=100, lookback=60, target=30, stride=90, n_features=3):
        self.device = torch.device(device)
        self.seed, self.S, self.T, self.n, self.stride, self.F = int(seed), n_stocks, lookback, target, stride, n_features
        self.lib = _lib.load()
        self._bufs = {}
        self.launches = 0

    def _alloc(self, B):
        if B not in self._bufs:
            N, d = B * self.S, self.device
            e = lambda *s: torch.empty(*s, dtype=torch.float32, device=d)
            self._bufs[B] = DeviceBatch(e(self.T, N, self.F), e(N, self.n), e(B, self.n), e(B, 2), e(N), B, self.S)
        return self._bufs[B]

    def generate(self, window_idx=None, first_window=0, n_windows=None, out=None, checked=False) -> DeviceBatch:
        if window_idx is not None:
            window_idx = window_idx.to(self.device, torch.int64).contiguous()
            n_windows = window_idx.numel()
        b = out if out is not None else self._alloc(int(n_windows))
        stream = torch.cuda.current_stream(self.device).cuda_stream
        check(self.lib.b200_gen_batch(self.seed, ptr(window_idx), int(first_window), int(n_windows), self.S, self.T, self.n,
                                      self.stride, self.F, ptr(b.x), ptr(b.y_stock), ptr(b.y_market), ptr(b.factor),
                                      ptr(b.inv_psi), stream), "b200_gen_batch")
        self.launches += 1
        return b

    def coefficients(self):
        a = torch.empty(self.S, dtype=torch.float32, device=self.device)
        b = torch.empty_like(a)
        stream = torch.cuda.current_stream(self.device).cuda_stream
        check(self.lib.b200_gen_coefficients(self.seed, self.S, ptr(a), ptr(b), stream), "b200_gen_coefficients")
        return a, b


class SeriesBatchSource:
    """Batches cut out of a device-resident series (stocks [S][n_samples], market [n_samples]) by b200_series_batch:
    the reference's prepare_data body (src/data.py:193-214) evaluated per batch on the device."""

    def __init__(self, device, stocks, market, lookback=60, target=30, stride=90, n_features=3, prediction_task=True,
                 alphas=None, betas=None):
        self.device = torch.device(device)
        self.stocks = stocks.to(self.device, torch.float32).contiguous()
        self.market = market.to(self.device, torch.float32).contiguous()
        if self.stocks.dim() != 2 or self.market.dim() != 1 or self.stocks.shape[1] != self.market.shape[0]:
            raise ValueError(f"series shapes do not match: stocks {tuple(stocks.shape)}, market {tuple(market.shape)}")
        self.S, self.n_samples = int(self.stocks.shape[0]), int(self.stocks.shape[1])
        self.T, self.n, self.stride, self.F = int(lookback), int(target), int(stride), int(n_features)
        self.prediction = bool(prediction_task)
        self.alphas = None if alphas is None else alphas.to(self.device, torch.float32)
        self.betas = None if betas is None else betas.to(self.device, torch.float32)
        self.lib = _lib.load()
        self._bufs = {}
        self.launches = 0

    @property
    def n_windows(self):
        total = self.T + self.n if self.prediction else self.T
        return (self.n_samples - total) // self.stride + 1           # unfold(size, step), src/common.py:101

    def _alloc(self, B):
        if B not in self._bufs:
            N, d = B * self.S, self.device
            e = lambda *s: torch.empty(*s, dtype=torch.float32, device=d)
            self._bufs[B] = DeviceBatch(e(self.T, N, self.F), e(N, self.n), e(B, self.n), e(B, 2), e(N), B, self.S, e(N), e(N))
        return self._bufs[B]

    def generate(self, window_idx=None, first_window=0, n_windows=None, out=None, checked=False) -> DeviceBatch:
        if window_idx is not None:
            window_idx = window_idx.to(torch.int64)
            if not checked and window_idx.numel() and (int(window_idx.min()) < 0 or int(window_idx.max()) >= self.n_windows):
                raise IndexError(f"window index outside [0, {self.n_windows})")
            window_idx = window_idx.to(self.device).contiguous()
            n_windows = window_idx.numel()
        b = out if out is not None else self._alloc(int(n_windows))
        stream = torch.cuda.current_stream(self.device).cuda_stream
        check(self.lib.b200_series_batch(ptr(self.stocks), ptr(self.market), self.n_samples, ptr(window_idx), int(first_window),
                                         int(n_windows), self.S, self.T, self.n, self.stride, self.F, int(self.prediction),
                                         ptr(b.x), ptr(b.y_stock), ptr(b.y_market), ptr(b.factor), ptr(b.inv_psi),
                                         ptr(b.ols_alpha), ptr(b.ols_beta), stream), "b200_series_batch")
        self.launches += 1
        return b

    def coefficients(self):
        """True per-stock (alpha, beta) when the series came with them (alphas.pt / betas.pt), else None: the reference
        then stores the per-window OLS coefficients (DeviceBatch.ols_alpha / ols_beta), src/data.py:205-207."""
        if self.alphas is None or self.betas is None:
            return None
        return self.alphas, self.betas


def generator_series(device, seed, n_stocks, n_samples, first_day=0):
    """The generator's own series on the device (what SyntheticLogReturns.generate returns, src/data.py:49-59)."""
    lib = _lib.load()
    dev = torch.device(device)
    stocks = torch.empty(n_stocks, n_samples, dtype=torch.float32, device=dev)
    market = torch.empty(n_samples, dtype=torch.float32, device=dev)
    stream = torch.cuda.current_stream(dev).cuda_stream
    check(lib.b200_gen_series(int(seed), int(first_day), int(n_samples), int(n_stocks), ptr(stocks), ptr(market), stream),
          "b200_gen_series")
    return stocks, market


def load_fama_french_25(raw_dir):
    """Host-side mirror of FamaFrench25Portfolios.load (src/data.py:86-123): the 25 size/book-to-market portfolios and
    the market factor from Kenneth French's daily CSVs -> excess log returns in percent, (p25 [25][n], mkt [n])."""
    import numpy as np
    import pandas as pd
    raw_dir = Path(raw_dir)
    n_samples, skip_old = 26129, 3125                                   # src/data.py:68-70
    ff3 = pd.read_csv(raw_dir / 'F-F_Research_Data_Factors_daily.csv', header=0, index_col=0,
                      names=['DATE', 'Mkt-RF', 'SMB', 'HML', 'RF'], usecols=['DATE', 'Mkt-RF', 'RF'],
                      skiprows=4 + skip_old, nrows=n_samples - skip_old)
    p25 = pd.read_csv(raw_dir / '25_Portfolios_5x5_Daily.csv', header=0, index_col=0,
                      names=['DATE'] + [f'P{i}' for i in range(25)], skiprows=18 + skip_old, nrows=n_samples - skip_old)
    MKT = torch.tensor(ff3['Mkt-RF'].values.astype(np.float32))
    RF = torch.tensor(ff3['RF'].values.astype(np.float32))
    P25 = torch.tensor(p25.values.astype(np.float32)).T - RF
    missing = (P25 == -99.99).any(dim=0) | (P25 == -999).any(dim=0)
    P25, MKT = P25[:, ~missing], MKT[~missing]
    log100 = torch.log(torch.tensor(100.0))
    return 100 * (torch.log(P25 + 100) - log100), 100 * (torch.log(MKT + 100) - log100)


class HostBatchFeeder:
    """Feeds batches that live in pinned HOST memory in the reference DataLoader's layout
    ([X (B,S,T,F), y (B,S,n,4), factor (B,2), inv_psi (B,S)], src/data.py:236-238) to the device path: the
    host->device copy and the layout conversion of batch k+1 run on a copy stream while batch k trains."""

    def __init__(self, device):
        self.device = torch.device(device)
        self.copy_stream = torch.cuda.Stream(device=self.device)

    def _start(self, host_batch):
        with torch.cuda.stream(self.copy_stream):
            dev = [t.to(self.device, non_blocking=True) for t in host_batch]
            b = DeviceBatch.from_reference(dev, self.device)
            ev = torch.cuda.Event()
            ev.record(self.copy_stream)
        return b, ev

    def _finish(self, started):
        b, ev = started
        cur = torch.cuda.current_stream(self.device)
        cur.wait_event(ev)
        for t in (b.x, b.y_stock, b.y_market, b.factor, b.inv_psi):
            t.record_stream(cur)
        return b

    def feed(self, host_batches):
        it = iter(host_batches)
        try:
            nxt = self._start(next(it))
        except StopIteration:
            return
        for hb in it:
            cur, nxt = nxt, self._start(hb)
            yield self._finish(cur)
        yield self._finish(nxt)


class _Loader:
    """Iterable of DeviceBatch over a window range; mirrors DataLoader(batch_size, shuffle)."""

    def __init__(self, source, lo, hi, batch_size, shuffle, generator=None, rank=0, world=1, drop_uneven=False):
        self.source, self.lo, self.hi, self.bs = source, lo, hi, int(batch_size)
        self.shuffle, self.gen, self.rank, self.world = shuffle, generator, rank, world

    def __len__(self):
        return (self.hi - self.lo + self.bs - 1) // self.bs

    def __iter__(self):
        n = self.hi - self.lo
        order = torch.randperm(n, generator=self.gen) if self.shuffle else torch.arange(n)
        # the epoch's order goes to the device ONCE (a per-batch upload of the indices is a blocking copy that makes the host
        # wait for the previous step: 2.2 of 2.4 ms per step at batch_size 1); batches are slices of it
        order = (order + self.lo).to(self.source.device)
        for i in range(0, n, self.bs):
            idx = order[i:i + self.bs]
            if self.world > 1:                        # each rank takes its contiguous share of the global batch
                per = idx.numel() // self.world
                if per == 0:
                    continue
                idx = idx[self.rank * per:(self.rank + 1) * per]
            yield self.source.generate(window_idx=idx, checked=True)     # lo..hi lie inside the series by construction


class _StreamLoader:
    """Endless fresh windows: batch k of rank r = windows first + k*B + [r*B/W, (r+1)*B/W)."""

    def __init__(self, source, first, batch_size, steps, rank=0, world=1):
        self.source, self.first, self.bs, self.steps, self.rank, self.world = source, first, int(batch_size), steps, rank, world

    def __len__(self):
        return self.steps

    def __iter__(self):
        per = self.bs // self.world
        for k in range(self.steps):
            yield self.source.generate(first_window=self.first + k * self.bs + self.rank * per, n_windows=per)


class FinancialLstmDataModule:
    """Constructor arguments as src/data.py:142-150; extra keyword arguments configure the device sources.

    Source selection (prepare_data): `data_dir/stocks.pt` + `market.pt` present -> file-backed series (the reference's
    own on-disk format, src/data.py:193-196); else `data_dir.name == "real"` -> built from `data_dir/../real_raw/*.csv`
    like train.py:22-26 (error if the CSVs are missing, the reference's TODO at train.py:20); else the generator."""

    def __init__(self, data_dir: Path = Path("data/synthetic"), lookback_window: int = 60, target_window: int = 20,
                 stride: int = 80, prediction_task: bool = True, interaction_only: bool = True, batch_size: int = 1,
                 *, seed: int = 1234, n_stocks: int = 100, n_samples: int = 1_000_000, device="cuda:0",
                 streaming: bool = False, steps_per_epoch: int | None = None, rank: int = 0, world_size: int = 1,
                 use_files: bool | None = None):
        if not prediction_task and target_window > lookback_window:
            raise ValueError('target window must be <= lookback window for reconstruction task')
        self.data_dir = Path(data_dir)
        self.hparams = dict(lookback_window=lookback_window, target_window=target_window, stride=stride,
                            prediction_task=prediction_task, interaction_only=interaction_only, batch_size=batch_size)
        self.seed, self.n_stocks, self.n_samples = seed, n_stocks, n_samples
        self.device, self.streaming, self.steps_per_epoch = device, streaming, steps_per_epoch
        self.rank, self.world_size = rank, world_size
        self.use_files = use_files
        self.n_features = 3 if interaction_only else 5
        self.source = None
        self.source_kind = None
        self._series = None
        self.n_windows = self.train_end = self.val_end = None
        self._shuffle_gen = torch.Generator().manual_seed(seed)

    def _files_present(self):
        return (self.data_dir / 'stocks.pt').exists() and (self.data_dir / 'market.pt').exists()

    def prepare_data(self):
        hp = self.hparams
        use_files = self._files_present() if self.use_files is None else self.use_files
        if not use_files and self.data_dir.name == "real":
            raw = self.data_dir.parent / "real_raw"
            if not (raw / '25_Portfolios_5x5_Daily.csv').exists():
                raise FileNotFoundError(f"datamodule=real needs the Fama-French CSVs in {raw} (reference train.py:20 leaves the "
                                        "download as a TODO); they are not shipped with this repository")
            p25, mkt = load_fama_french_25(raw)                      # train.py:22-26
            self.data_dir.mkdir(parents=True, exist_ok=True)
            torch.save(p25, self.data_dir / 'stocks.pt')
            torch.save(mkt, self.data_dir / 'market.pt')
            use_files = True
        if use_files:
            load = lambda name: torch.load(self.data_dir / name) if (self.data_dir / name).exists() else None
            stocks, market = load('stocks.pt'), load('market.pt')
            if stocks is None or market is None:
                raise FileNotFoundError(f"{self.data_dir} holds no stocks.pt / market.pt")
            self._series = (stocks.float(), market.float(), load('alphas.pt'), load('betas.pt'))
            self.n_stocks, self.n_samples = int(stocks.shape[0]), int(stocks.shape[1])
            self.source_kind = "files"
        else:
            self.source_kind = "generator" if hp["prediction_task"] else "generator-series"
        total = hp["lookback_window"] + (hp["target_window"] if hp["prediction_task"] else 0)
        self.n_windows = (self.n_samples - total) // hp["stride"] + 1           # unfold(size, step), src/common.py:101
        self.train_end, self.val_end = int(0.7 * self.n_windows), int(0.9 * self.n_windows)   # src/data.py:225-227

    def setup(self, stage: str | None = None):
        if self.n_windows is None:
            self.prepare_data()
        if self.source is not None:
            return
        hp = self.hparams
        if self.source_kind == "generator":
            self.source = DeviceBatchSource(self.device, self.seed, self.n_stocks, hp["lookback_window"],
                                            hp["target_window"], hp["stride"], self.n_features)
            return
        if self.source_kind == "files":
            stocks, market, alphas, betas = self._series
        else:                                   # reconstruction split of the generator's series
            stocks, market = generator_series(self.device, self.seed, self.n_stocks, self.n_samples)
            alphas, betas = DeviceBatchSource(self.device, self.seed, self.n_stocks).coefficients()
        self.source = SeriesBatchSource(self.device, stocks, market, hp["lookback_window"], hp["target_window"], hp["stride"],
                                        self.n_features, hp["prediction_task"], alphas, betas)
        self._series = None

    def train_dataloader(self):
        bs = self.hparams["batch_size"]
        if self.streaming:
            if self.source_kind != "generator":
                raise ValueError("streaming=True needs the generator source (a finite series has no fresh windows)")
            steps = self.steps_per_epoch or max(1, self.train_end // max(1, bs))
            # stream from beyond the test split so that evaluation windows are never trained on
            return _StreamLoader(self.source, self.n_windows, bs, steps, self.rank, self.world_size)
        return _Loader(self.source, 0, self.train_end, bs, True, self._shuffle_gen, self.rank, self.world_size)

    def val_dataloader(self, batch_size=1):
        return _Loader(self.source, self.train_end, self.val_end, batch_size, False)

    def test_dataloader(self, batch_size=1):
        return _Loader(self.source, self.val_end, self.n_windows, batch_size, False)

    def teardown(self, stage: str):
        pass
