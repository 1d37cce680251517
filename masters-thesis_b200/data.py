"""Device-resident mirror of the reference datamodule (src/data.py:133-250).

Same constructor arguments and the same prepare_data()/setup()/..._dataloader() surface, but no
dataset file is materialised: every window is a pure function of (seed, window index) evaluated by
the `b200_gen_batch` kernel (generator spec: oracle/lstm_oracle.py), so a batch is produced on the
device from a list of window indices.  The reference's 70/20/10 contiguous split of the
(n_samples - 90)//stride + 1 windows (src/data.py:221-234) and its shuffled train loader
(src/data.py:236-238) are kept; `streaming=True` drops the epoch structure and walks fresh windows
(needed once batch_size exceeds the 7,777 training windows of the default 1,000,000-day series).
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
        """Inverse of to_reference(): accepts the list of four tensors the reference's loaders yield."""
        X, y, factor, inv_psi = [t.to(device=device, dtype=torch.float32) for t in batch]
        B, S, T, F = X.shape
        n = y.shape[2]
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

    def generate(self, window_idx=None, first_window=0, n_windows=None, out=None) -> DeviceBatch:
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
        order = order + self.lo
        for i in range(0, n, self.bs):
            idx = order[i:i + self.bs]
            if self.world > 1:                        # each rank takes its contiguous share of the global batch
                per = idx.numel() // self.world
                if per == 0:
                    continue
                idx = idx[self.rank * per:(self.rank + 1) * per]
            yield self.source.generate(window_idx=idx)


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
    """Constructor arguments as src/data.py:142-150; extra keyword arguments configure the device generator."""

    def __init__(self, data_dir: Path = Path("data/synthetic"), lookback_window: int = 60, target_window: int = 20,
                 stride: int = 80, prediction_task: bool = True, interaction_only: bool = True, batch_size: int = 1,
                 *, seed: int = 1234, n_stocks: int = 100, n_samples: int = 1_000_000, device="cuda:0",
                 streaming: bool = False, steps_per_epoch: int | None = None, rank: int = 0, world_size: int = 1):
        if not prediction_task and target_window > lookback_window:
            raise ValueError('target window must be <= lookback window for reconstruction task')
        if not prediction_task:
            raise NotImplementedError("prediction_task=false (reconstruction) is outside the B200 hot path (SURVEY.md f4)")
        self.data_dir = Path(data_dir)
        self.hparams = dict(lookback_window=lookback_window, target_window=target_window, stride=stride,
                            prediction_task=prediction_task, interaction_only=interaction_only, batch_size=batch_size)
        self.seed, self.n_stocks, self.n_samples = seed, n_stocks, n_samples
        self.device, self.streaming, self.steps_per_epoch = device, streaming, steps_per_epoch
        self.rank, self.world_size = rank, world_size
        self.n_features = 3 if interaction_only else 5
        self.source = None
        self.n_windows = self.train_end = self.val_end = None
        self._shuffle_gen = torch.Generator().manual_seed(seed)

    def prepare_data(self):
        hp = self.hparams
        total = hp["lookback_window"] + hp["target_window"]
        self.n_windows = (self.n_samples - total) // hp["stride"] + 1           # unfold(size, step), src/common.py:101
        self.train_end, self.val_end = int(0.7 * self.n_windows), int(0.9 * self.n_windows)   # src/data.py:225-227

    def setup(self, stage: str | None = None):
        if self.n_windows is None:
            self.prepare_data()
        if self.source is None:
            hp = self.hparams
            self.source = DeviceBatchSource(self.device, self.seed, self.n_stocks, hp["lookback_window"],
                                            hp["target_window"], hp["stride"], self.n_features)

    def train_dataloader(self):
        bs = self.hparams["batch_size"]
        if self.streaming:
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
