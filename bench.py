"""Headline benchmark: training throughput (samples = windows per second) of the per-step training
loop on the synthetic datamodule, B200-native arm and reference (CPU oracle port) arm.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--engine ffma|tc_precise|tc_fast]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

One step = gen_batch -> LSTM forward -> objective -> BPTT -> global-norm clip -> Adam = ONE optimiser step.
  N = 1: BASELINE.json configs[1] -- defaults small/mse/fast at datamodule.batch_size 4096 on one B200.
  N > 1: BASELINE.json configs[3] -- the same model data parallel at GLOBAL batch 65,536 windows (65,536 / N per GPU,
         run as micro-batches of 4096 windows inside one optimiser step; gradients all-reduced over NCCL in buckets
         while the last micro-batch's lower layers are still in backward).  A second, weak-scaling number at 4096
         windows per GPU is reported under "weak_4096_per_gpu" (--batch overrides the per-GPU batch of the headline).
Prints ONE JSON line (rank 0).  See DESIGN.md "Measurement" for every field.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

T, N_TARGET, S = 60, 30, 100
MODELS = {"small": (64, 2, 0.2), "medium": (64, 4, 0.3), "large": (64, 8, 0.3), "wide": (256, 8, 0.3)}
CLIP, LR = 5.0, 1e-4          # configs/trainer/fast.yaml, configs/model/small.yaml
GLOBAL_DP_BATCH = 65536       # BASELINE.json configs[3]
TRAFFIC_FILE = ROOT / "profiles" / "ncu_traffic.json"


def flops_per_sequence(F, H, L):
    """Algorithmic training flops per sequence: 3*T*sum_l 8H(I_l+H) - T*8H*I_0 + 6*2H."""
    tot = sum(8 * H * ((F if l == 0 else H) + H) for l in range(L))
    return 3 * T * tot - T * 8 * H * F + 6 * 2 * H


def peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        d = json.loads(p.read_text())
        return d.get("hbm_gbs", 6650.0), d.get("bf16_tflops", 1590.0), d.get("bf16_tflops_sustained", 1400.0), "measured"
    return 6650.0, 1590.0, 1400.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index=0):
        self.rows, self.proc, self.idx = [], None, gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-i", str(self.idx), "-lms", "100"], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[1])); mx.append(float(r[2]))
                for n, v in zip(names, r[4:8]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except (ValueError, IndexError):
                continue
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


# --------------------------------------------------------------------------------------------- reference arm
def oracle_steps_per_second(model_name, loss, batch_windows, steps, warmup, threads, seed=1234, compiled=False):
    """Times the oracle port of the reference's training step on the host cores; returns (windows/s, s/step)."""
    import torch
    sys.path.insert(0, str(ROOT / "oracle"))
    import lstm_oracle as O
    torch.set_num_threads(threads)
    torch.set_float32_matmul_precision("highest")
    H, L, p = MODELS[model_name]
    torch.manual_seed(seed)
    m = O.OracleLstm(3, H, L, p)
    if compiled:                        # the reference calls model.compile() (train.py:137): torch.compile of the module, in place
        m.compile()
    opt = O.make_optimizer(m, LR)
    batches = [O.generate_windows(range(i * batch_windows, (i + 1) * batch_windows), seed=seed) for i in range(2)]
    for i in range(warmup):
        O.train_step(m, opt, batches[i % 2], loss, CLIP)
    t0 = time.perf_counter()
    for i in range(steps):
        O.train_step(m, opt, batches[i % 2], loss, CLIP)
    dt = time.perf_counter() - t0
    return batch_windows * steps / dt, dt / steps


def cpu_variants(args, cores):
    """The other CPU lines BASELINE.md section 3 plans: the reference default batch_size 1 (configs/datamodule/synthetic.yaml:7)
    and the model.compile() variant (train.py:137).  Each is a bounded sample; a variant that cannot run reports why."""
    out = {}
    for name, kw, bw, steps, warm in (("eager_batch1", {}, 1, 100, 10), ("compiled_batch64", {"compiled": True}, args.cpu_batch, 4, 3)):
        try:
            v, sec = oracle_steps_per_second(args.model, args.loss, bw, steps, warm, cores, **kw)
            out[name] = {"value": v, "unit": "samples/s", "s_per_step": sec, "windows_per_step": bw, "steps": steps}
        except Exception as exc:            # e.g. no host compiler for torch.compile
            out[name] = {"unavailable": f"{type(exc).__name__}: {str(exc)[:160]}"}
    return out


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    bw = args.cpu_batch
    v, sec = oracle_steps_per_second(args.model, args.loss, bw, args.steps, args.warmup, cores)
    world = int(os.environ.get("WORLD_SIZE", "1"))
    batch = per_gpu_batch(args, world)
    sample = (f"each step = one optimisation step of the oracle port on {bw} windows ({bw * S} sequences), "
              f"a 1/{max(1, batch // bw)} sample of the {batch}-window batch; torch CPU fp32 eager, {cores} threads")
    cfg = workload_config(args, world, batch)
    out = {"impl": "reference", "metric": "train samples/sec", "value": v, "unit": "samples/s", "n_gpus": args.gpus,
           "steps": args.steps, "warmup": args.warmup, "ms_per_step": sec * 1e3, "higher_is_better": True,
           "scaling": "weak" if world == 1 or args.batch else "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
           "config": cfg, "cpu_baseline": {"value": v, "unit": "samples/s", "cores": cores, "kind": "port", "sample": sample},
           "e2e": {"value": v, "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(out), flush=True)


def per_gpu_batch(args, world):
    if args.batch:
        return args.batch
    return 4096 if world == 1 else GLOBAL_DP_BATCH // world


def workload_config(args, world, batch):
    H, L, p = MODELS[args.model]
    micro = min(batch, args.micro)
    if world == 1 or args.batch:
        what = f"at datamodule.batch_size={batch} per GPU"
    else:
        what = (f"data parallel at GLOBAL batch {batch * world} windows (BASELINE.json configs[3]): {batch} windows per GPU and optimiser "
                f"step")
    return {"workload": f"configs/config.yaml defaults (datamodule=synthetic model={args.model} loss={args.loss} trainer=fast "
                        f"clip {CLIP}) {what}: {batch * S} sequences x {T} steps per GPU and step"
                        + (f" in {(batch + micro - 1) // micro} micro-batches of {micro} windows" if batch > micro else "")
                        + f", H={H} L={L} dropout={p}",
            "global_batch_windows": batch * world, "sequences_per_step": batch * S * world, "lookback": T,
            "micro_batch_windows": micro, "engine": args.engine,
            "reference_arm": f"`--impl reference` (the CPU oracle port on all host cores) times one optimisation step on a {args.cpu_batch}-window "
                             "SAMPLE of this batch per timed step; samples/s is per window, so the two arms compare like for like",
            "l2": "per-step working set (activation stash, >= 25 GB at 4096 windows) exceeds the 126 MB L2",
            "parallelism": f"dp{world}"}


# --------------------------------------------------------------------------------------------- b200 arm
class Run:
    """One model + datamodule + trainer on this rank."""

    def __init__(self, args, model_name, batch, rank, world, dev, engine=None, micro=None):
        import torch
        import masters_thesis_b200 as B
        H, L, p = MODELS[model_name]
        torch.manual_seed(1234)
        cls = {"mse": B.FinancialLstmMse, "nll": B.FinancialLstmNll, "combined": B.FinancialLstmCombined}[args.loss]
        self.model = cls(3, H, L, p, LR)
        self.dm = B.FinancialLstmDataModule(lookback_window=T, target_window=N_TARGET, stride=90, batch_size=batch, seed=1234,
                                            device=dev, streaming=True, steps_per_epoch=1, rank=rank, world_size=world)
        self.dm.prepare_data(); self.dm.setup("fit")
        self.tr = B.B200Trainer(max_epochs=1, gradient_clip_val=CLIP, engine=engine or args.engine, verbose=False,
                                micro_batch_windows=micro or args.micro)
        self.tr._attach(self.model, self.dm)
        self.tr._init_distributed()
        self.e = self.model.engine
        self.batch, self.rank, self.world = batch, rank, world
        self.micro = self.tr.micro_windows(self.e, batch)
        self.e.ensure_capacity(min(batch, self.micro))
        self.first = self.dm.n_windows

    def step(self, k):
        b = self.dm.source.generate(first_window=self.first + (k * self.world + self.rank) * self.batch, n_windows=self.batch,
                                    micro=self.micro)
        return self.tr.training_step(self.model, b, LR)

    def free(self):
        import torch
        self.tr.close()
        self.model.engine = None
        self.e = self.dm = self.tr = self.model = None
        import gc
        gc.collect()
        torch.cuda.empty_cache()


def timed_steps(run, steps, warmup, barrier, dist, dev, lib, timing=True):
    """W warm-up steps, then exactly K steps between barriers; CUDA events, max over ranks.  Returns (ms, kernel table, losses)."""
    import torch
    from masters_thesis_b200 import _lib
    for k in range(warmup):
        run.step(k)
    barrier()
    if timing:
        lib.b200_timing_enable(1)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    ev0.record()
    for k in range(steps):
        losses = run.step(warmup + k)
    ev1.record()
    barrier()
    ms = ev0.elapsed_time(ev1)
    kt = _lib.timing_report() if timing else {}
    if timing:
        lib.b200_timing_enable(0)
    t_ms = torch.tensor([ms], dtype=torch.float64, device=dev)
    if run.world > 1:
        dist.all_reduce(t_ms, op=dist.ReduceOp.MAX)
    return float(t_ms), kt, losses


def lstm_roofline(kt, steps, seqs, H, L, tf_sus, src):
    F_seq = flops_per_sequence(3, H, L)
    mine = {k: v for k, v in kt.items() if not k.startswith("nccl")}
    dom = max(mine.items(), key=lambda kv: kv[1][1]) if mine else (None, (0, 0.0))
    lstm_ms = sum(v[1] for k, v in mine.items() if k.startswith(("lstm_", "wide_", "w2_")))
    per_step = lstm_ms / steps if steps else 0.0
    achieved = seqs * F_seq / (per_step * 1e-3) / 1e12 if per_step else 0.0
    return {"bound": "tensor", "kernel": "LSTM forward+backward kernels (" + ", ".join(sorted(k for k in mine if k.startswith(("lstm_", "wide_", "w2_")))) + ")",
            "dominant_kernel": dom[0], "achieved": achieved, "peak": tf_sus, "unit": "TFLOP/s", "frac": achieved / tf_sus,
            "peak_source": f"{src} bf16 sustained (MEASURED_PEAKS.json)", "algorithmic_flops_per_step": seqs * F_seq,
            "lstm_ms_per_step": per_step,
            "kernel_ms_per_step": {k: v[1] / steps for k, v in sorted(kt.items())},
            "kernel_launches_per_step": {k: v[0] / steps for k, v in sorted(kt.items())}}


def torch_gpu_line(args, batch, dev):
    """Informational: the stock PyTorch path of the reference on the SAME GPU and batch -- torch.nn.LSTM (cuDNN) + Linear heads +
    mse + clip_grad_norm_ + torch.optim.Adam (src/model.py:88-109, 150-154, 204-209; train.py:13 'medium' matmul precision).
    BASELINE.md section 4 names it the kernel to beat.  Not the reference arm (that is the CPU path), not a parity target."""
    import torch
    import masters_thesis_b200 as B
    H, L, p = MODELS[args.model]

    class Stock(torch.nn.Module):          # the reference's module structure (src/model.py:88-109) from stock torch parts
        def __init__(self):
            super().__init__()
            self.lstm = torch.nn.LSTM(3, H, L, dropout=p if L > 1 else 0, batch_first=True)
            self.alpha_head, self.beta_head = torch.nn.Linear(H, 1), torch.nn.Linear(H, 1)

        def forward(self, x):
            last = self.lstm(x)[0][:, -1, :]
            return self.alpha_head(last), self.beta_head(last)

    def step(m, opt, ref):
        X, y = ref[0].flatten(0, 1), ref[1].flatten(0, 1)
        a, b = m(X)
        loss = torch.mean((a + b * y[:, :, 1] - y[:, :, 0]) ** 2)          # src/model.py:199-206 (mse objective)
        opt.zero_grad()
        loss.backward()
        torch.nn.utils.clip_grad_norm_(m.parameters(), CLIP)
        opt.step()

    out = {"what": "torch.nn.LSTM (cuDNN) + heads + mse + clip_grad_norm_ + torch.optim.Adam, fp32 with "
                   "set_float32_matmul_precision('medium') as train.py:13, mse objective, device-resident batch", "informational": True}
    prev = torch.get_float32_matmul_precision()
    try:
        torch.set_float32_matmul_precision("medium")
        torch.manual_seed(1234)
        m = Stock().to(dev).train()
        opt = torch.optim.Adam(m.parameters(), lr=LR, weight_decay=1e-5)
        src = B.data.DeviceBatchSource(dev, 1234)
        bw = batch
        while True:
            try:
                ref = src.generate(first_window=20000, n_windows=bw).to_reference()
                for _ in range(3):
                    step(m, opt, ref)
                torch.cuda.synchronize()
                break
            except torch.OutOfMemoryError:
                ref = None
                torch.cuda.empty_cache()
                bw //= 2
                if bw < 64:
                    raise
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n = 5
        ev0.record()
        for _ in range(n):
            step(m, opt, ref)
        ev1.record()
        torch.cuda.synchronize()
        ms = ev0.elapsed_time(ev1) / n
        out.update({"value": bw / (ms / 1e3), "unit": "samples/s", "ms_per_step": ms, "batch_windows": bw, "steps": n})
    except Exception as exc:
        out["unavailable"] = f"{type(exc).__name__}: {str(exc)[:200]}"
    finally:
        torch.set_float32_matmul_precision(prev)
        torch.cuda.empty_cache()
    return out


def run_b200(args):
    import torch
    import torch.distributed as dist
    import masters_thesis_b200 as B
    from masters_thesis_b200 import _lib

    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dev = torch.device(f"cuda:{local}")
    torch.cuda.set_device(dev)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    H, L, p = MODELS[args.model]
    lib = _lib.load()
    for o in args.opt:
        k, _, v = o.partition("=")
        _lib.check(lib.b200_set_option(k.encode(), int(v)), f"b200_set_option({o})")
    batch = per_gpu_batch(args, world)
    strong = world > 1 and not args.batch

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    run = Run(args, args.model, batch, rank, world, dev)
    e, dm, tr, model = run.e, run.dm, run.tr, run.model

    # ---- device-resident throughput -------------------------------------------------------------
    clocks = ClockSampler(local)
    for k in range(args.warmup):
        run.step(k)
    barrier()
    if rank == 0:
        clocks.start()
    l0 = e.launches + dm.source.launches
    ms, kt, losses = timed_steps(run, args.steps, 0, barrier, dist, dev, lib)
    # the library's own per-name launch counts of the timed region (TimedLaunch registry; NCCL's kernel is not ours);
    # the engine's host-side counter is the fallback
    launches = int(sum(v[0] for k, v in kt.items() if not k.startswith("nccl"))) if kt else e.launches + dm.source.launches - l0
    clk = clocks.stop() if rank == 0 else None
    value = batch * world * args.steps / (ms / 1e3)
    final_loss = float(losses[2])

    # ---- end to end: reference-format batches in pinned host memory -> H2D -> step -> loss D2H ----
    e2e = None
    if not args.no_e2e:
        # two distinct batches in pinned host memory, alternated
        refs = []
        for i in range(2):
            parts = []
            for lo in range(0, batch, 4096):            # built in pieces: to_reference() of the whole batch would double device memory
                nb = min(4096, batch - lo)
                parts.append([t.cpu() for t in dm.source.generate(first_window=run.first + i * batch + lo, n_windows=nb).to_reference()])
            refs.append([torch.cat([q[j] for q in parts]).pin_memory() for j in range(4)])
        h2d = sum(t.numel() * 4 for t in refs[0])
        e2e_steps = max(2, args.e2e_steps if args.e2e_steps else args.steps)
        feeder = B.HostBatchFeeder(dev)

        def run_e2e(n):
            # every step's loss is read back to the host (pinned buffer + event), one step behind the enqueue front
            box, vals = B.LossMailbox(dev), []
            for b in feeder.feed(refs[i % 2] for i in range(n)):
                vals += box.post(tr.training_step(model, b, LR))
            vals += box.drain()
            assert len(vals) == n
            return vals[-1][2]

        run_e2e(3)
        barrier()
        t0 = time.perf_counter()
        run_e2e(e2e_steps)
        barrier()
        dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        e2e = {"value": batch * world * e2e_steps / float(dt), "unit": "samples/s", "h2d_bytes_per_step": h2d,
               "d2h_bytes_per_step": 16, "steps": e2e_steps,
               "path": "pinned host [X,y,factor,inv_psi] (reference DataLoader layout) -> HostBatchFeeder (cudaMemcpyAsync + b200_collate_reference on a copy stream, one batch ahead) -> B200Trainer.training_step -> LossMailbox (every step's [mse, nll, total, mae] copied to pinned host memory and read by the host one step behind the enqueue front)"}
        del refs, feeder

    # ---- second number under data parallelism: weak scaling at 4096 windows per GPU --------------------------
    weak = None
    if strong:
        e = dm = tr = model = None
        run.free()
        run = Run(args, args.model, 4096, rank, world, dev)
        e, dm, tr, model = run.e, run.dm, run.tr, run.model
        wms, wkt, _ = timed_steps(run, args.steps, args.warmup, barrier, dist, dev, lib)
        weak = {"value": 4096 * world * args.steps / (wms / 1e3), "unit": "samples/s", "ms_per_step": wms / args.steps,
                "scaling": "weak", "global_batch_windows": 4096 * world,
                "kernel_ms_per_step": {k: v[1] / args.steps for k, v in sorted(wkt.items())}}
    del e, dm, tr, model, losses
    run.free()

    # ---- rank 0 only from here: other configurations and baselines -------------------------------------------
    extra_cfgs, tgpu = [], None
    hbm, tf_burst, tf_sus, src = peaks()
    if world == 1 and rank == 0:
        if not args.no_torch_gpu:
            tgpu = torch_gpu_line(args, batch, dev)
        if not args.no_wide and args.model != "wide":
            # BASELINE.json configs[2]: widened model (4x hidden width, 8 layers), bf16 GEMMs / fp32 accumulate, same batch of
            # 4096 windows per optimiser step, run as micro-batches sized to the device's memory
            try:
                wa = argparse.Namespace(**vars(args))
                wa.loss, wa.engine = args.loss, "tc_fast"
                wrun = Run(wa, "wide", args.wide_batch, 0, 1, dev, engine="tc_fast")
                wsteps = 2
                # ~17,000 launches per optimiser step at this size: the per-launch CUDA events of the kernel table would sit in the
                # launch path, so the throughput is timed WITHOUT them and the per-kernel table comes from a separate step
                wms, _, wl = timed_steps(wrun, wsteps, 1, barrier, dist, dev, lib, timing=False)
                _, wkt, _ = timed_steps(wrun, 1, 0, barrier, dist, dev, lib, timing=True)
                wh, wl_, wp = MODELS["wide"]
                roof = lstm_roofline(wkt, 1, args.wide_batch * S, wh, wl_, tf_sus, src)
                # roofline of this line: algorithmic flops over the WHOLE step time (the LSTM kernels are > 99 % of it).  The per-kernel
                # CUDA-event table of the ~15,000 short launches also counts the gaps between an event and its kernel, so its sum can
                # exceed the step; it is kept for the shares only.
                roof["kernel_table_frac"] = roof["frac"]
                roof["achieved"] = args.wide_batch * S * flops_per_sequence(3, wh, wl_) / (wms / wsteps * 1e-3) / 1e12
                roof["frac"] = roof["achieved"] / tf_sus
                extra_cfgs.append({"name": "BASELINE.json configs[2]: widened model hidden 256 x 8 layers, bf16 GEMMs / fp32 accumulate (engine tc_fast), 1xB200",
                                   "metric": "train samples/sec", "value": args.wide_batch * wsteps / (wms / 1e3), "unit": "samples/s",
                                   "ms_per_step": wms / wsteps, "steps": wsteps, "warmup": 1, "dtype": "bf16",
                                   "config": {"model": "wide", "batch_windows": args.wide_batch, "micro_batch_windows": wrun.micro,
                                              "loss": args.loss, "dropout": wp},
                                   "final_loss": float(wl[2]), "roofline": roof})
                wrun.free()
            except Exception as exc:
                extra_cfgs.append({"name": "BASELINE.json configs[2] widened model", "unavailable": f"{type(exc).__name__}: {str(exc)[:200]}"})
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rank != 0:
        return
    # ---- roofline of the LSTM kernels (live CUDA-event durations from the timed region) -------
    roofline = lstm_roofline(kt, args.steps, batch * S, H, L, tf_sus, src)
    traffic = None
    if TRAFFIC_FILE.exists():   # DRAM bytes per launch of the dominant kernel from the committed ncu --set full capture of this workload
        t = json.loads(TRAFFIC_FILE.read_text())
        wl = t.get("workload", {})
        if (wl.get("model"), wl.get("engine"), wl.get("batch")) == (args.model, args.engine, min(batch, args.micro)):
            traffic = t["bytes_per_launch"].get(roofline["dominant_kernel"])
    roofline["traffic"] = traffic
    roofline["traffic_source"] = (f"committed ncu --set full capture ({TRAFFIC_FILE.relative_to(ROOT)}: dram__bytes_read.sum + dram__bytes_write.sum "
                                  "per launch of the dominant kernel on this workload); NOT measured in this run") if traffic else None
    # ---- CPU baseline: bounded sample on this box's host cores -----------------------------------
    cpu = None
    if not args.no_cpu:
        cores = os.cpu_count() or 1
        v, sec = oracle_steps_per_second(args.model, args.loss, args.cpu_batch, args.cpu_steps, 2, cores)
        cpu = {"value": v, "unit": "samples/s", "cores": cores, "kind": "port",
               "sample": f"{args.cpu_steps} optimisation steps of the oracle port at {args.cpu_batch} windows/step "
                         f"(1/{max(1, batch // args.cpu_batch)} of the batch), torch CPU fp32 eager, {sec:.2f} s/step"}
        if world == 1 and not args.no_cpu_variants:
            cpu["variants"] = cpu_variants(args, cores)
    out = {"metric": "train samples/sec", "value": value, "unit": "samples/s", "n_gpus": world, "steps": args.steps,
           "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
           "scaling": "strong" if strong else "weak",
           "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": workload_config(args, world, batch),
           "sequences_per_sec": value * S, "final_loss": final_loss, "clocks": clk, "e2e": e2e, "gpu_launches": launches,
           "roofline": roofline, "cpu_baseline": cpu}
    if weak:
        out["weak_4096_per_gpu"] = weak
    if tgpu:
        out["torch_gpu"] = tgpu
    if extra_cfgs:
        out["configs"] = extra_cfgs
    print(json.dumps(out), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--engine", default="tc_precise", help="ffma | tc_precise | tc_fast")
    ap.add_argument("--model", default="small", choices=sorted(MODELS))
    ap.add_argument("--loss", default="mse", choices=["mse", "nll", "combined"])
    ap.add_argument("--batch", type=int, default=0, help="windows per GPU per optimiser step (default: 4096 at N = 1, 65536 / N at N > 1)")
    ap.add_argument("--micro", type=int, default=4096, help="windows per forward/backward pass (micro-batch cap)")
    ap.add_argument("--cpu-batch", type=int, default=64, help="windows per step of the CPU sample")
    ap.add_argument("--cpu-steps", type=int, default=6)
    ap.add_argument("--e2e-steps", type=int, default=0, help="timed end-to-end steps (default: --steps)")
    ap.add_argument("--wide-batch", type=int, default=4096, help="windows per optimiser step of the configs[2] line")
    ap.add_argument("--dropout", type=float, default=None, help="override the model's inter-layer dropout (diagnostics)")
    ap.add_argument("--opt", action="append", default=[], help="library tuning knob name=value (b200_set_option), repeatable")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-cpu-variants", action="store_true")
    ap.add_argument("--no-torch-gpu", action="store_true")
    ap.add_argument("--no-wide", action="store_true")
    ap.add_argument("--kernels-only", action="store_true", help="shorthand for --no-e2e --no-cpu --no-torch-gpu --no-wide")
    args = ap.parse_args()
    if args.kernels_only:
        args.no_e2e = args.no_cpu = args.no_torch_gpu = args.no_wide = True
    if args.dropout is not None:
        h, l, _ = MODELS[args.model]
        MODELS[args.model] = (h, l, args.dropout)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
