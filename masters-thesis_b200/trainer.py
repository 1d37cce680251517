"""B200Trainer: the object `train.py` builds in place of lightning.Trainer when `trainer=b200`.

It restates Lightning's fit/test loops around the reference modules (train.py:169-198): epochs of
shuffled training batches -> one fused device step each (forward, objective, BPTT, global-norm clip,
Adam), validation with batch size 1 semantics every `check_val_every_n_epoch` epochs, ReduceLROnPlateau
on loss/total/val (src/model.py:156-172), best/last checkpoints (train.py:151-161) and the test loop
(src/model.py:119-147).  Under torchrun (WORLD_SIZE > 1) it is data parallel: each rank takes its share
of every epoch like Lightning's DistributedSampler (`datamodule.batch_size` is per rank, global batch = W x
batch_size) and gradients are summed (pre-scaled by 1/W) with bucketed all-reduces (b200_dp_allreduce: NCCL from
inside the library) issued per layer on a communication stream while the lower layers' backward kernels are still
running.  A local batch whose activations do not fit the device is processed in micro-batches inside ONE optimiser
step (engine.forward_backward).
"""
from __future__ import annotations

import math
import os
import time
from pathlib import Path

import torch

from .grad_sync import GradBucketer


class LossMailbox:
    """Per-step device -> host read of the step's [mse, nll, total, mae] without draining the GPU queue: step k's
    values are copied to pinned host memory on the compute stream and read by the host while step k + 1 is already
    enqueued (a `.item()` after every step would leave the GPU idle for one host launch sequence per step)."""

    def __init__(self, device, depth=2):
        self.device = torch.device(device)
        self.slots = [torch.empty(4, dtype=torch.float32).pin_memory() for _ in range(depth)]
        self.events = [torch.cuda.Event() for _ in range(depth)]
        self.pending = []            # slot indices in flight, oldest first
        self.n = 0

    def post(self, losses):
        """Queue the read of one step's losses; returns the values of steps that fell out of the window."""
        done = []
        if len(self.pending) == len(self.slots):
            done.append(self._take())
        i = self.n % len(self.slots)
        self.slots[i].copy_(losses, non_blocking=True)
        self.events[i].record(torch.cuda.current_stream(self.device))
        self.pending.append(i)
        self.n += 1
        return done

    def _take(self):
        i = self.pending.pop(0)
        self.events[i].synchronize()
        return self.slots[i].tolist()

    def drain(self):
        out = []
        while self.pending:
            out.append(self._take())
        return out


class PlateauSchedule:
    """torch.optim.lr_scheduler.ReduceLROnPlateau(mode='min', factor, patience) with torch's defaults
    (relative threshold 1e-4, cooldown 0, min_lr 0, eps 1e-8) -- src/model.py:156-161."""

    def __init__(self, lr, factor=0.5, patience=2, threshold=1e-4, eps=1e-8):
        self.lr, self.factor, self.patience, self.threshold, self.eps = lr, factor, patience, threshold, eps
        self.best, self.num_bad = math.inf, 0

    def step(self, metric):
        if metric < self.best * (1.0 - self.threshold):
            self.best, self.num_bad = metric, 0
        else:
            self.num_bad += 1
        if self.num_bad > self.patience:
            new_lr = self.lr * self.factor
            if self.lr - new_lr > self.eps:
                self.lr = new_lr
            self.num_bad = 0
        return self.lr


class _CheckpointState:
    """The two attributes of ModelCheckpoint that train.py reads (train.py:193-194, 210, 220)."""

    def __init__(self, dirpath=None):
        self.dirpath = dirpath
        self.best_model_path = ""
        self.best_model_score = None
        self.last_model_path = ""


def _precision_to_mode(precision, engine):
    if engine and engine != "auto":
        return engine
    return {"32-true": "tc_precise", "32": "tc_precise", "bf16-mixed": "tc_fast", "bf16": "tc_fast"}.get(str(precision), "tc_precise")


class B200Trainer:
    def __init__(self, max_epochs=4, gradient_clip_val=None, precision="32-true", enable_progress_bar=False,
                 enable_model_summary=True, logger=None, callbacks=None, check_val_every_n_epoch=1, *,
                 devices=1, max_steps=-1, seed=1234, bucket_mb=0.05, engine="auto", eval_batch_windows=512,
                 checkpoint_dir=None, log_dir=None, limit_val_windows=None, limit_test_windows=None, verbose=True,
                 micro_batch_windows=4096):
        self.max_epochs, self.clip = int(max_epochs), gradient_clip_val
        self.precision, self.check_val_every_n_epoch = precision, int(check_val_every_n_epoch)
        self.max_steps, self.seed, self.bucket_mb = int(max_steps), int(seed), float(bucket_mb)
        self.mode = _precision_to_mode(precision, engine)
        self.eval_batch_windows = int(eval_batch_windows)
        self.micro_batch_windows = int(micro_batch_windows) if micro_batch_windows else 0
        self.limit_val_windows, self.limit_test_windows = limit_val_windows, limit_test_windows
        self.verbose = verbose and int(os.environ.get("RANK", "0")) == 0
        self.checkpoint_callback = _CheckpointState(checkpoint_dir)
        self.log_dir = log_dir
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local_rank = int(os.environ.get("LOCAL_RANK", "0"))
        self.device = torch.device(f"cuda:{self.local_rank}")
        self.global_step, self.current_epoch = 0, 0
        self.history = []            # one dict of logged metrics per epoch
        self.step_losses = []        # device tensors [mse, nll, total, mae] per step when record_steps
        self.record_steps = False
        self._tb = None
        self._sync = None
        self._comm = None
        self._micro_cache = {}
        # optional callable(engine, global_step) invoked after the gradients of a step are complete (all-reduced under
        # data parallelism) and before clip + Adam consume them: Lightning's on_before_optimizer_step hook
        self.on_before_optimizer_step = None

    # ------------------------------------------------------------------ plumbing
    def _init_distributed(self):
        """torch.distributed is the rendezvous (it carries the NCCL unique id and the per-epoch scalar reductions); the
        gradient exchange itself goes through the library's own communicator (b200_dp_init / b200_dp_allreduce)."""
        if self.world > 1 and self._sync is None:
            import ctypes
            import torch.distributed as dist
            from . import _lib
            if not dist.is_initialized():
                dist.init_process_group("nccl", device_id=self.device)
            lib = _lib.load()
            torch.cuda.set_device(self.device)
            uid = ctypes.create_string_buffer(128)
            if self.rank == 0:
                _lib.check(lib.b200_dp_unique_id(uid), "b200_dp_unique_id")
            box = [uid.raw]
            dist.broadcast_object_list(box, src=0)
            comm = ctypes.c_void_p()
            _lib.check(lib.b200_dp_init(box[0], self.world, self.rank, ctypes.byref(comm)), "b200_dp_init")
            self._comm = comm
            self._sync = GradBucketer(self.world, self.bucket_mb, device=self.device, comm=comm, lib=lib)

    def close(self):
        """Destroy the library communicator (optional; the process exit does the same)."""
        if self._comm is not None:
            from . import _lib
            _lib.load().b200_dp_destroy(self._comm)
            self._comm, self._sync = None, None

    def micro_windows(self, engine, batch_windows):
        """Windows per micro-batch: the configured cap, lowered when the device's free memory cannot hold the activation
        stash and workspaces of that many windows (engine buffers already allocated count as free)."""
        key = (id(engine), int(batch_windows))
        if key in self._micro_cache:                       # cudaMemGetInfo per step would sit in the launch path
            return self._micro_cache[key]
        cap = self.micro_batch_windows or batch_windows
        free, _ = torch.cuda.mem_get_info(self.device)
        held = sum(getattr(engine, n).numel() * getattr(engine, n).element_size()
                   for n in ("stash_h", "stash_c", "lstm_ws") if hasattr(engine, n))
        fit = int(0.85 * (free + held) / engine.bytes_per_window())
        m = max(1, min(cap, batch_windows, fit))
        m = m if m >= batch_windows else max(1, m // 64 * 64 if m >= 64 else m)
        self._micro_cache[key] = m
        return m

    def _writer(self):
        if self._tb is None and self.log_dir and self.rank == 0:
            try:
                from torch.utils.tensorboard import SummaryWriter
                self._tb = SummaryWriter(self.log_dir)
            except Exception:            # tensorboard missing: scalars stay in self.history
                self._tb = False
        return self._tb or None

    def _log(self, metrics, epoch):
        self.history.append(dict(metrics, epoch=epoch))
        w = self._writer()
        if w:
            for k, v in metrics.items():
                w.add_scalar(k, v, epoch)

    def _attach(self, model, dm):
        torch.cuda.set_device(self.device)
        if model.engine is None:
            hp = dm.hparams
            model.attach(self.device, self.mode, hp["lookback_window"], hp["target_window"], dm.n_stocks)
        return model.engine

    def dropout_seed(self, step):
        return (self.seed * 1000003 + step) & 0xFFFFFFFFFFFFFFFF

    # ------------------------------------------------------------------ one step
    def training_step(self, model, batch, lr):
        """One optimiser step on this rank's batch (a DeviceBatch, or a ChunkedBatch of micro-batches)."""
        e = model.engine
        if getattr(batch, "chunks", None) is None:       # a whole batch handed in by the caller (host-fed path): split here
            m = self.micro_windows(e, batch.n_windows)
            if batch.n_windows > m:
                from .data import ChunkedBatch
                batch = ChunkedBatch.split(batch, m)
        sync = self._sync
        ready = None
        if sync is not None:
            sync.begin(e.grads, e.P)
            ready = sync.layer_ready                      # (offset, length, last): heads slice first, then layers top-down
        losses = e.forward_backward(batch, model.loss_kind, model.hparams.get("mse_weight", 100.0),
                                    self.dropout_seed(self.global_step), 1.0 / self.world, ready)
        if sync is not None:
            sync.wait()
        if self.on_before_optimizer_step is not None:
            self.on_before_optimizer_step(e, self.global_step)
        e.optimizer_step(lr, self.clip if self.clip else 0.0, model.hparams["weight_decay"])
        self.global_step += 1
        return losses

    # ------------------------------------------------------------------ loops
    @torch.no_grad()
    def evaluate(self, model, dm, split="val"):
        """Batch-1 semantics over the split, processed `eval_batch_windows` windows per launch."""
        e = self._attach(model, dm)
        lo, hi = (dm.train_end, dm.val_end) if split == "val" else (dm.val_end, dm.n_windows)
        limit = self.limit_val_windows if split == "val" else self.limit_test_windows
        if limit:
            hi = min(hi, lo + int(limit))
        # data parallel: every rank evaluates a contiguous share of the split; the sums are all-reduced (every window
        # is its own batch, so the split's means do not depend on how it is cut)
        sharded = self.world > 1 and self._sync is not None
        if sharded:
            per = (hi - lo + self.world - 1) // self.world
            lo, hi = min(hi, lo + self.rank * per), min(hi, lo + (self.rank + 1) * per)
        acc = torch.zeros(5, dtype=torch.float64, device=self.device)
        done = 0
        kind = model.loss_kind if split == "val" else "combined"
        nb_max = min(self.eval_batch_windows, self.micro_windows(e, self.eval_batch_windows))
        while done < hi - lo:
            nb = min(nb_max, hi - lo - done)
            batch = dm.source.generate(first_window=lo + done, n_windows=nb)
            acc[:4] += e.eval_windows(batch, kind, model.hparams.get("mse_weight", 100.0)).double() * nb
            done += nb
        acc[4] = done
        if sharded:
            import torch.distributed as dist
            dist.all_reduce(acc)
        done = int(acc[4].item())
        mse, nll, total, mae = (acc[:4] / max(done, 1)).tolist()
        return {"mse": mse, "nll": nll, "total": total, "mae": mae, "windows": done}

    def fit(self, model, datamodule=None):
        dm = datamodule
        dm.rank, dm.world_size, dm.device = self.rank, self.world, self.device
        dm.prepare_data()
        dm.setup("fit")
        e = self._attach(model, dm)
        self._init_distributed()
        if getattr(dm, "micro_batch_windows", None) is None:
            dm.micro_batch_windows = self.micro_windows(e, int(dm.hparams["batch_size"]))
        sched = PlateauSchedule(model.hparams["learning_rate"])
        ck = self.checkpoint_callback
        kind = model.loss_kind
        stop = False
        for epoch in range(self.max_epochs):
            self.current_epoch = epoch
            t0 = time.time()
            loader = dm.train_dataloader()
            acc = torch.zeros(4, dtype=torch.float64, device=self.device)
            nb = 0
            for batch in loader:
                losses = self.training_step(model, batch, sched.lr)
                acc += losses.double()
                nb += 1
                if self.record_steps:
                    self.step_losses.append(losses.clone())
                if 0 < self.max_steps <= self.global_step:
                    stop = True
                    break
            if self.world > 1:
                import torch.distributed as dist
                dist.all_reduce(acc)
                acc /= self.world
            mse, nll, total, _ = (acc / max(nb, 1)).tolist()
            metrics = {"loss/total/train": total, "lr-Adam": sched.lr}
            if kind in ("mse", "combined"):
                metrics["loss/mse/train"] = mse
            if kind in ("nll", "combined"):
                metrics["loss/nll/train"] = nll
            # validation, scheduler and checkpoints run on Lightning's cadence only (multiples of check_val_every_n_epoch,
            # also when max_steps ends the epoch early): best score, lr history and checkpoint names then match the reference
            if (epoch + 1) % self.check_val_every_n_epoch == 0:
                v = self.evaluate(model, dm, "val")
                metrics["loss/total/val"] = v["total"]
                if kind in ("mse", "combined"):
                    metrics["loss/mse/val"] = v["mse"]
                if kind in ("nll", "combined"):
                    metrics["loss/nll/val"] = v["nll"]
                sched.step(v["total"])
                if ck.best_model_score is None or v["total"] < float(ck.best_model_score):
                    ck.best_model_score = torch.tensor(v["total"])
                    ck.best_model_path = self._save(model, f"best_epoch={epoch:02d}_val={v['total']:.4f}.ckpt", epoch)
                ck.last_model_path = self._save(model, "last.ckpt", epoch)
            metrics["epoch_seconds"] = time.time() - t0
            self._log(metrics, epoch)
            if self.verbose:
                print(f"[b200] epoch {epoch}: " + ", ".join(f"{k}={v:.6g}" for k, v in metrics.items()), flush=True)
            if stop:
                break
        if ck.best_model_score is None:
            ck.best_model_score = torch.tensor(float("nan"))
        return self

    def test(self, model, datamodule=None):
        dm = datamodule
        dm.prepare_data()
        dm.setup("test")
        self._attach(model, dm)
        model.on_test_start()
        t = self.evaluate(model, dm, "test")
        model.test_results['test_loss']['mae'] = t["mae"]       # src/model.py:142-147
        model.test_results['test_loss']['nll'] = t["nll"]
        if self.verbose:
            print(f"[b200] test: mae={t['mae']:.6f} nll={t['nll']:.6f} over {t['windows']} windows", flush=True)
        return [{"test/mae": t["mae"], "test/nll": t["nll"]}]

    def _save(self, model, filename, epoch):
        """Lightning-layout checkpoint (state_dict + hyper_parameters), consumable by load_from_checkpoint
        (test.py:178).  Rank 0 only."""
        ck = self.checkpoint_callback
        if not ck.dirpath or self.rank != 0:
            return ""
        d = Path(ck.dirpath)
        d.mkdir(parents=True, exist_ok=True)
        path = d / filename
        torch.save({"state_dict": model.state_dict(), "hyper_parameters": dict(model.hparams), "epoch": epoch,
                    "global_step": self.global_step, "pytorch-lightning_version": "2.6.0"}, path)
        return str(path)
