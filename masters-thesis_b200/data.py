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


class ChunkedBatch:
    """One optimiser step's local batch as micro-batches: the activation stash of a whole batch can exceed device memory
    (25 GB per 4096 windows for the small model), so the LSTM runs micro-batch by micro-batch while the step keeps the
    reference's semantics -- one loss, one clip, one Adam update per `datamodule.batch_size` windows
    (src/data.py:236-238, src/model.py:204-209).  The targets of the WHOLE batch live in one set of tensors (the joint
    NLL of src/model.py:237-247 couples every row of the batch); each chunk's y_stock / y_market / factor / inv_psi
    are views into them, x is per chunk (time-major, so a row range of it is not contiguous)."""

    def __init__(self, chunks, y_stock, y_market, factor, inv_psi, n_stocks):
        self.chunks = chunks
        self.y_stock, self.y_market, self.factor, self.inv_psi = y_stock, y_market, factor, inv_psi
        self.n_windows = sum(c.n_windows for c in chunks)
        self.n_stocks = n_stocks

    @staticmethod
    def split(batch: "DeviceBatch", micro: int):
        """Micro-batches of an existing DeviceBatch (copies the x rows of each chunk; targets are views)."""
        B, S = batch.n_windows, batch.n_stocks
        chunks = []
        for lo in range(0, B, micro):
            hi = min(B, lo + micro)
            chunks.append(DeviceBatch(batch.x[:, lo * S:hi * S].contiguous(), batch.y_stock[lo * S:hi * S], batch.y_market[lo:hi],
                                      batch.factor[lo:hi], batch.inv_psi[lo * S:hi * S], hi - lo, S))
        return ChunkedBatch(chunks, batch.y_stock, batch.y_market, batch.factor, batch.inv_psi, S)


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

    def generate(self, window_idx=None, first_window=0, n_windows=None, out=None, checked=False, micro=None):
        """One batch; with `micro` (windows) smaller than the batch, a ChunkedBatch of micro-batches."""
        if window_idx is not None:
            window_idx = window_idx.to(self.device, torch.int64).contiguous()
            n_windows = window_idx.numel()
        n_windows = int(n_windows)
        if micro and n_windows > int(micro):
            return self._generate_chunked(window_idx, int(first_window), n_windows, int(micro))
        b = out if out is not None else self._alloc(n_windows)
        self._launch(window_idx, int(first_window), n_windows, b.x, b.y_stock, b.y_market, b.factor, b.inv_psi)
        return b

    def _launch(self, window_idx, first_window, n_windows, x, y_stock, y_market, factor, inv_psi):
        stream = torch.cuda.current_stream(self.device).cuda_stream
        check(self.lib.b200_gen_batch(self.seed, ptr(window_idx), first_window, n_windows, self.S, self.T, self.n,
                                      self.stride, self.F, ptr(x), ptr(y_stock), ptr(y_market), ptr(factor),
                                      ptr(inv_psi), stream), "b200_gen_batch")
        self.launches += 1

    def _generate_chunked(self, window_idx, first_window, n_windows, micro):
        key = ("chunked", n_windows, micro)
        if key not in self._bufs:
            N, d = n_windows * self.S, self.device
            e = lambda *s: torch.empty(*s, dtype=torch.float32, device=d)
            ys, ym, fa, ip = e(N, self.n), e(n_windows, self.n), e(n_windows, 2), e(N)
            chunks = []
            for lo in range(0, n_windows, micro):
                hi = min(n_windows, lo + micro)
                chunks.append(DeviceBatch(e(self.T, (hi - lo) * self.S, self.F), ys[lo * self.S:hi * self.S], ym[lo:hi], fa[lo:hi],
                                          ip[lo * self.S:hi * self.S], hi - lo, self.S))
            self._bufs[key] = ChunkedBatch(chunks, ys, ym, fa, ip, self.S)
        cb = self._bufs[key]
        lo = 0
        for c in cb.chunks:
            idx = None if window_idx is None else window_idx[lo:lo + c.n_windows]
            self._launch(idx, first_window + lo, c.n_windows, c.x, c.y_stock, c.y_market, c.factor, c.inv_psi)
            lo += c.n_windows
        return cb

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

    def generate(self, window_idx=None, first_window=0, n_windows=None, out=None, checked=False, micro=None):
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
        if micro and int(n_windows) > int(micro):
            return ChunkedBatch.split(b, int(micro))
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


def rank_share(order, rank, world):
    """The windows of one epoch that rank `rank` of `world` sees: what torch's DistributedSampler (which Lightning
    installs under DDP, train.py:169) does to the epoch's order -- pad by wrapping around to a multiple of the world
    size, then every world-th index starting at the rank.  world == 1 returns the order unchanged."""
    if world <= 1:
        return order
    n = order.numel()
    total = (n + world - 1) // world * world
    if total > n:
        order = torch.cat([order, order[: total - n]])
    return order[rank:total:world]


class _Loader:
    """Iterable of DeviceBatch over a window range; mirrors DataLoader(batch_size, shuffle).  Under data parallelism
    `batch_size` is PER RANK, as with Lightning DDP on the reference (global batch = world x batch_size): every rank
    shuffles with the same generator and takes its DistributedSampler share (rank_share) of the epoch."""

    def __init__(self, source, lo, hi, batch_size, shuffle, generator=None, rank=0, world=1, micro=None):
        self.source, self.lo, self.hi, self.bs = source, lo, hi, int(batch_size)
        self.shuffle, self.gen, self.rank, self.world, self.micro = shuffle, generator, rank, max(1, world), micro
        if self.bs < 1:
            raise ValueError(f"batch_size must be >= 1, got {batch_size}")

    def _per_rank(self):
        return (self.hi - self.lo + self.world - 1) // self.world

    def __len__(self):
        return (self._per_rank() + self.bs - 1) // self.bs

    def __iter__(self):
        n = self.hi - self.lo
        order = torch.randperm(n, generator=self.gen) if self.shuffle else torch.arange(n)
        order = rank_share(order + self.lo, self.rank, self.world)
        # the epoch's order goes to the device ONCE (a per-batch upload of the indices is a blocking copy that makes the host
        # wait for the previous step: 2.2 of 2.4 ms per step at batch_size 1); batches are slices of it
        order = order.to(self.source.device)
        for i in range(0, order.numel(), self.bs):
            # lo..hi lie inside the series by construction
            yield self.source.generate(window_idx=order[i:i + self.bs], checked=True, micro=self.micro)


class _StreamLoader:
    """Endless fresh windows, `batch_size` per rank: step k of rank r = windows first + (k*W + r)*B .. +B."""

    def __init__(self, source, first, batch_size, steps, rank=0, world=1, micro=None):
        self.source, self.first, self.bs, self.steps = source, first, int(batch_size), steps
        self.rank, self.world, self.micro = rank, max(1, world), micro

    def __len__(self):
        return self.steps

    def __iter__(self):
        for k in range(self.steps):
            yield self.source.generate(first_window=self.first + (k * self.world + self.rank) * self.bs, n_windows=self.bs,
                                       micro=self.micro)


class FinancialLstmDataModule:
    """Constructor arguments as src/data.py:142-150; extra keyword arguments configure the device sources.

    Source selection (prepare_data): `data_dir/stocks.pt` + `market.pt` present -> file-backed series (the reference's
    own on-disk format, src/data.py:193-196); else `data_dir.name == "real"` -> built from `data_dir/../real_raw/*.csv`
    like train.py:22-26 (error if the CSVs are missing, the reference's TODO at train.py:20); else the generator."""

    def __init__(self, data_dir: Path = Path("data/synthetic"), lookback_window: int = 60, target_window: int = 20,
                 stride: int = 80, prediction_task: bool = True, interaction_only: bool = True, batch_size: int = 1,
                 *, seed: int = 1234, n_stocks: int = 100, n_samples: int = 1_000_000, device="cuda:0",
                 streaming: bool = False, steps_per_epoch: int | None = None, rank: int = 0, world_size: int = 1,
                 use_files: bool | None = None, micro_batch_windows: int | None = None):
        if not prediction_task and target_window > lookback_window:
            raise ValueError('target window must be <= lookback window for reconstruction task')
        self.data_dir = Path(data_dir)
        self.hparams = dict(lookback_window=lookback_window, target_window=target_window, stride=stride,
                            prediction_task=prediction_task, interaction_only=interaction_only, batch_size=batch_size)
        self.seed, self.n_stocks, self.n_samples = seed, n_stocks, n_samples
        self.device, self.streaming, self.steps_per_epoch = device, streaming, steps_per_epoch
        self.rank, self.world_size = rank, world_size
        self.use_files = use_files
        self.micro_batch_windows = micro_batch_windows      # set by the trainer from the device's free memory when None
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
            steps = self.steps_per_epoch or max(1, self.train_end // max(1, bs * max(1, self.world_size)))
            # stream from beyond the test split so that evaluation windows are never trained on
            return _StreamLoader(self.source, self.n_windows, bs, steps, self.rank, self.world_size, self.micro_batch_windows)
        return _Loader(self.source, 0, self.train_end, bs, True, self._shuffle_gen, self.rank, self.world_size,
                       self.micro_batch_windows)

    def val_dataloader(self, batch_size=1):
        return _Loader(self.source, self.train_end, self.val_end, batch_size, False)

    def test_dataloader(self, batch_size=1):
        return _Loader(self.source, self.val_end, self.n_windows, batch_size, False)

    def teardown(self, stage: str):
        pass
