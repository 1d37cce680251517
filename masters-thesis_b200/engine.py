"""Device-side step engine: owns the flat parameter/optimiser vectors, the activation stash and the
workspaces (torch tensors are the memory owner), and drives the C-ABI kernels on torch's current
stream.  One instance per process (= per GPU).

The sequence of one optimisation step mirrors Lightning's automatic optimisation of the reference
(train.py:169-187, src/model.py:204-209): forward -> loss -> (zero_grad) backward -> clip -> Adam.
"""
from __future__ import annotations

import torch

from . import _lib
from ._lib import LOSS_KINDS, MODES, check, ptr


class LstmEngine:
    def __init__(self, n_features=3, hidden=64, layers=2, dropout=0.0, device="cuda:0", mode="ffma",
                 lookback=60, target=30, n_stocks=100):
        if not torch.cuda.is_available():
            raise _lib.B200Error("LstmEngine needs a CUDA device; there is no CPU fallback")
        self.lib = _lib.load()
        self.F, self.H, self.L, self.p = int(n_features), int(hidden), int(layers), float(dropout)
        self.T, self.n, self.S = int(lookback), int(target), int(n_stocks)
        self.mode = MODES[mode] if isinstance(mode, str) else int(mode)
        # hidden sizes every kernel on the path supports: 64 (persistent engines, all modes) and 128 / 256 (wide tcgen05
        # engine x heads kernels); checked here so that an unsupported width fails before any launch, not mid-step
        if not (self.H == 64 or (self.H in (128, 256) and self.mode != MODES["ffma"])):
            raise _lib.B200Error(f"hidden_size={self.H} is not supported with engine mode {mode}: the fp32 CUDA-core engine "
                                 "runs hidden 64 only, the tcgen05 engines hidden 64, 128 and 256")
        self.device = torch.device(device)
        self.P = int(self.lib.b200_param_count(self.F, self.H, self.L))
        self.head_off = int(self.lib.b200_param_offset(self.F, self.H, self.L, self.L))
        self.P_pad = (self.P + 3) // 4 * 4
        z = lambda: torch.zeros(self.P_pad, dtype=torch.float32, device=self.device)
        self.params, self.grads, self.exp_avg, self.exp_avg_sq = z(), z(), z(), z()
        self.losses = torch.zeros(4, dtype=torch.float32, device=self.device)
        self.sqnorm = torch.zeros(1, dtype=torch.float32, device=self.device)
        self.sq_ws = torch.empty(int(self.lib.b200_sqnorm_workspace_bytes(self.P_pad)), dtype=torch.uint8, device=self.device)
        self.opt_step = 0
        self._cap = 0          # capacity in windows
        self.launches = 0      # kernel launches issued (bench.py's gpu_launches claim)

    # ------------------------------------------------------------------ buffers
    def ensure_capacity(self, n_windows: int):
        if n_windows <= self._cap:
            return
        N = n_windows * self.S
        dev = self.device
        self.stash_h = torch.empty(self.L * self.T * N * self.H, dtype=torch.float32, device=dev)
        self.stash_c = torch.empty_like(self.stash_h)
        self.dhT = torch.empty(N * self.H, dtype=torch.float32, device=dev)
        self.alpha = torch.empty(N, dtype=torch.float32, device=dev)
        self.beta = torch.empty(N, dtype=torch.float32, device=dev)
        ws = int(self.lib.b200_lstm_workspace_bytes(self.mode, N, self.T, self.F, self.H, self.L))
        self.lstm_ws = torch.empty(max(ws, 256), dtype=torch.uint8, device=dev)
        hws = int(self.lib.b200_heads_workspace_bytes(n_windows, self.S, self.n, self.H))
        self.heads_ws = torch.empty(max(hws, 256), dtype=torch.uint8, device=dev)
        self._cap = n_windows

    def stash_bytes(self, n_windows: int) -> int:
        return 2 * self.L * self.T * n_windows * self.S * self.H * 4

    @property
    def stream(self):
        return torch.cuda.current_stream(self.device).cuda_stream

    def load_params(self, flat: torch.Tensor):
        assert flat.numel() == self.P, (flat.numel(), self.P)
        self.params[: self.P].copy_(flat.to(self.device, torch.float32))

    def reset_optimizer(self):
        self.exp_avg.zero_(); self.exp_avg_sq.zero_(); self.grads.zero_(); self.opt_step = 0

    # ------------------------------------------------------------------ kernels
    def forward(self, batch, training=False, dropout_seed=0):
        """LSTM forward of every layer; fills the stash.  Replaces self.lstm(...) (src/model.py:104)."""
        self.ensure_capacity(batch.n_windows)
        N = batch.n_windows * self.S
        check(self.lib.b200_lstm_fwd(self.mode, ptr(self.params), self.F, self.H, self.L, ptr(batch.x), N, self.T,
                                     ptr(self.stash_h), ptr(self.stash_c), self.p, int(dropout_seed), int(bool(training)),
                                     ptr(self.lstm_ws), self.lstm_ws.numel(), self.stream), "b200_lstm_fwd")
        self.launches += self.L
        self._N = N

    def hidden_states(self, layer, n_windows):
        """fp32 [T][N][H] hidden states of one layer after forward() (lstm_out of src/model.py:104 for that layer)."""
        N = n_windows * self.S
        out = torch.empty(self.T, N, self.H, dtype=torch.float32, device=self.device)
        check(self.lib.b200_lstm_hidden_states(self.mode, int(layer), self.F, self.H, self.L, N, self.T, ptr(self.stash_h),
                                               ptr(self.lstm_ws), self.lstm_ws.numel(), ptr(out), self.stream), "b200_lstm_hidden_states")
        return out

    def final_hidden(self, n_windows):
        N = n_windows * self.S
        off = ((self.L - 1) * self.T + (self.T - 1)) * N * self.H
        return self.stash_h[off: off + N * self.H]

    def heads(self, batch, kind="mse", mse_weight=100.0, windows_per_group=None, grad_scale=1.0, want_grad=True,
              hT=None, dhT=None, head_grads=None, alpha=None, beta=None, heads_ws=None):
        """Heads + objective (+ gradients into dhT and the head slice of grads).  losses -> self.losses.
        `batch` supplies the targets (a DeviceBatch or a ChunkedBatch: its whole-batch tensors); the keyword buffers
        default to the engine's per-micro-batch ones."""
        B = batch.n_windows
        wpg = B if windows_per_group is None else int(windows_per_group)
        k = LOSS_KINDS[kind] if isinstance(kind, str) else int(kind)
        hT = self.final_hidden(B) if hT is None else hT
        head_params = self.params[self.head_off:]
        head_grads = self.grads[self.head_off:] if head_grads is None else head_grads
        dhT = self.dhT if dhT is None else dhT
        alpha, beta = (self.alpha if alpha is None else alpha), (self.beta if beta is None else beta)
        ws = self.heads_ws if heads_ws is None else heads_ws
        check(self.lib.b200_heads_loss_fwd_bwd(
            k, float(mse_weight), ptr(head_params), self.H, ptr(hT), ptr(batch.y_stock), ptr(batch.y_market),
            ptr(batch.factor), ptr(batch.inv_psi), B, self.S, self.n, wpg, float(grad_scale),
            ptr(alpha), ptr(beta), ptr(self.losses), ptr(dhT) if want_grad else None,
            ptr(head_grads) if want_grad else None, ptr(ws), ws.numel(), self.stream),
            "b200_heads_loss_fwd_bwd")
        self.launches += 5 if want_grad else 3
        return self.losses

    def backward_layer(self, batch, layer, training=True, dropout_seed=0, dhT=None, grads=None):
        N = batch.n_windows * self.S
        dhT = self.dhT if dhT is None else dhT
        grads = self.grads if grads is None else grads
        check(self.lib.b200_lstm_bwd_layer(self.mode, layer, ptr(self.params), self.F, self.H, self.L, ptr(batch.x), N,
                                           self.T, ptr(self.stash_h), ptr(self.stash_c), ptr(dhT), ptr(grads),
                                           self.p, int(dropout_seed), int(bool(training)), ptr(self.lstm_ws),
                                           self.lstm_ws.numel(), self.stream), "b200_lstm_bwd_layer")
        self.launches += 3

    def backward(self, batch, training=True, dropout_seed=0, on_layer_done=None):
        """BPTT, top layer first; on_layer_done(layer) is called after each layer's launches are queued."""
        for layer in range(self.L - 1, -1, -1):
            self.backward_layer(batch, layer, training, dropout_seed)
            if on_layer_done is not None:
                on_layer_done(layer)

    def layer_slice(self, layer):
        """(offset, length) in floats of a layer's gradient slice; layer == L gives the heads."""
        lo = int(self.lib.b200_param_offset(self.F, self.H, self.L, layer))
        hi = self.P if layer == self.L else int(self.lib.b200_param_offset(self.F, self.H, self.L, layer + 1))
        return lo, hi - lo

    def _accumulate(self, dst, src, count, scale=1.0):
        check(self.lib.b200_grad_accumulate(ptr(dst), ptr(src), int(count), float(scale), self.stream), "b200_grad_accumulate")
        self.launches += 1

    @staticmethod
    def chunk_seed(dropout_seed, chunk):
        """Dropout seed of micro-batch `chunk` of a step (the keep mask is a function of seed and row-in-launch, so every
        micro-batch needs its own stream of masks); chunk 0 keeps the step's seed."""
        return (int(dropout_seed) + 0x9E3779B97F4A7C15 * int(chunk)) & 0xFFFFFFFFFFFFFFFF

    def forward_backward(self, batch, kind, mse_weight=100.0, dropout_seed=0, grad_scale=1.0, on_slice_ready=None):
        """Loss and gradient of ONE optimiser step's local batch: self.grads <- grad_scale * d loss / d params, returns the
        device tensor [mse, nll, total, mae].  on_slice_ready(offset, length, last) is called as soon as every kernel
        writing a slice of self.grads is queued (heads first, then layers top-down): the data-parallel trainer hooks
        its bucket all-reduce there.

        A ChunkedBatch runs micro-batch by micro-batch (see data.ChunkedBatch):
          * mse is a mean over rows, so each micro-batch contributes its own loss and gradient with weight
            windows_in_chunk / windows_in_batch -- exact up to fp32 summation order;
          * nll / combined couple all rows of the batch through the joint Gaussian (src/model.py:237-247): pass 1 runs the
            forward of every micro-batch and collects the final hidden states, the heads kernels then see the WHOLE
            batch at once (one loss, dL/dh_T for every row), pass 2 re-runs each micro-batch's forward (same dropout
            masks) and back-propagates its rows."""
        ready = on_slice_ready or (lambda off, n, last: None)
        chunks = getattr(batch, "chunks", None)
        if chunks is None or len(chunks) == 1:
            b = batch if chunks is None else chunks[0]
            self.forward(b, training=True, dropout_seed=dropout_seed)
            losses = self.heads(b, kind, mse_weight, None, grad_scale, want_grad=True)
            ready(*self.layer_slice(self.L), False)
            for layer in range(self.L - 1, -1, -1):
                self.backward_layer(b, layer, True, dropout_seed)
                ready(*self.layer_slice(layer), layer == 0)
            return losses
        Btot, nchunk = batch.n_windows, len(chunks)
        self.ensure_capacity(max(c.n_windows for c in chunks))
        if not hasattr(self, "grads_chunk"):
            self.grads_chunk = torch.zeros_like(self.grads)
            self.losses_acc = torch.zeros_like(self.losses)
        joint = (kind if isinstance(kind, str) else {0: "mse", 1: "nll", 2: "combined"}[int(kind)]) != "mse"
        hl, hn = self.layer_slice(self.L)
        if joint:
            self._ensure_joint(Btot)
            row = 0
            for ci, c in enumerate(chunks):                         # pass 1: final hidden state of every row
                self.forward(c, training=True, dropout_seed=self.chunk_seed(dropout_seed, ci))
                n = c.n_windows * self.S * self.H
                self.hT_full[row:row + n].copy_(self.final_hidden(c.n_windows))
                row += n
            losses = self.heads(batch, kind, mse_weight, None, grad_scale, True, hT=self.hT_full, dhT=self.dhT_full,
                                alpha=self.alpha_full, beta=self.beta_full, heads_ws=self.heads_ws_full)
            ready(hl, hn, False)
        else:
            self.losses_acc.zero_()
        row = 0
        for ci, c in enumerate(chunks):
            first, last = ci == 0, ci == nchunk - 1
            seed = self.chunk_seed(dropout_seed, ci)
            gbuf = self.grads if first else self.grads_chunk       # the first micro-batch writes the step's gradient directly
            self.forward(c, training=True, dropout_seed=seed)
            n = c.n_windows * self.S * self.H
            if joint:
                dhT = self.dhT_full[row:row + n]
            else:
                w = c.n_windows / Btot
                self.heads(c, kind, mse_weight, None, grad_scale * w, True, head_grads=gbuf[self.head_off:])
                self._accumulate(self.losses_acc, self.losses, 4, w)
                if not first:
                    self._accumulate(self.grads[hl:], self.grads_chunk[hl:], hn)
                if last:
                    ready(hl, hn, False)
                dhT = self.dhT
            row += n
            for layer in range(self.L - 1, -1, -1):
                self.backward_layer(c, layer, True, seed, dhT=dhT, grads=gbuf)
                lo, ln = self.layer_slice(layer)
                if not first:
                    self._accumulate(self.grads[lo:], self.grads_chunk[lo:], ln)
                if last:
                    ready(lo, ln, layer == 0)
        if not joint:
            self.losses.copy_(self.losses_acc)
        return self.losses

    def _ensure_joint(self, n_windows):
        """Whole-batch buffers of the two-pass joint-NLL path: final hidden states, their gradient, alpha/beta, heads workspace."""
        if getattr(self, "_joint_cap", 0) >= n_windows:
            return
        N, dev = n_windows * self.S, self.device
        self.hT_full = torch.empty(N * self.H, dtype=torch.float32, device=dev)
        self.dhT_full = torch.empty_like(self.hT_full)
        self.alpha_full = torch.empty(N, dtype=torch.float32, device=dev)
        self.beta_full = torch.empty_like(self.alpha_full)
        hws = int(self.lib.b200_heads_workspace_bytes(n_windows, self.S, self.n, self.H))
        self.heads_ws_full = torch.empty(max(hws, 256), dtype=torch.uint8, device=dev)
        self._joint_cap = n_windows

    def bytes_per_window(self):
        """Device bytes the engine's per-micro-batch buffers need per window (stash, LSTM workspace, dhT ...): used by the
        trainer to size micro-batches for the free device memory."""
        probe = 1024
        ws = int(self.lib.b200_lstm_workspace_bytes(self.mode, probe * self.S, self.T, self.F, self.H, self.L))
        return self.stash_bytes(1) + ws // probe + self.S * (self.H + 2) * 4 + self.S * (self.T * self.F + self.n + 1) * 4

    def optimizer_step(self, lr, clip_val=0.0, weight_decay=1e-5, betas=(0.9, 0.999), eps=1e-8):
        """Global-norm clip + Adam on the flat vector (train.py:171, src/model.py:150-154)."""
        self.opt_step += 1
        use_clip = clip_val is not None and clip_val > 0
        if self.P_pad <= 131072:           # B200_CLIP_ADAM_FUSED_MAX: norm, clip and Adam in one launch
            check(self.lib.b200_clip_adam_step(ptr(self.params), ptr(self.grads), ptr(self.exp_avg), ptr(self.exp_avg_sq),
                                               self.P_pad, float(lr), float(betas[0]), float(betas[1]), float(eps),
                                               float(weight_decay), self.opt_step, float(clip_val) if use_clip else 0.0,
                                               ptr(self.sqnorm), self.stream), "b200_clip_adam_step")
            self.launches += 1
            return
        if use_clip:
            check(self.lib.b200_grad_sqnorm(ptr(self.grads), self.P_pad, ptr(self.sqnorm), ptr(self.sq_ws),
                                            self.sq_ws.numel(), self.stream), "b200_grad_sqnorm")
            self.launches += 2
        check(self.lib.b200_adam_step(ptr(self.params), ptr(self.grads), ptr(self.exp_avg), ptr(self.exp_avg_sq),
                                      self.P_pad, float(lr), float(betas[0]), float(betas[1]), float(eps),
                                      float(weight_decay), self.opt_step, float(clip_val) if use_clip else 0.0,
                                      ptr(self.sqnorm) if use_clip else None, 1, self.stream), "b200_adam_step")
        self.launches += 1

    def train_step(self, batch, kind, lr, clip_val, mse_weight=100.0, dropout_seed=0, grad_scale=1.0,
                   on_slice_ready=None, before_optimizer=None):
        """One optimisation step; returns the device tensor [mse, nll, total, mae] (no host sync)."""
        losses = self.forward_backward(batch, kind, mse_weight, dropout_seed, grad_scale, on_slice_ready)
        if before_optimizer is not None:
            before_optimizer()
        self.optimizer_step(lr, clip_val)
        return losses

    def eval_windows(self, batch, kind="combined", mse_weight=100.0):
        """Validation / test semantics: every window is its own batch (src/data.py:241-244); returns the
        device tensor [mean mse, mean per-window nll, mean total, mae]."""
        self.forward(batch, training=False)
        return self.heads(batch, kind, mse_weight, windows_per_group=1, want_grad=False)
