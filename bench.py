"""Headline benchmark: training throughput (samples = windows per second) of the per-step training
loop on the synthetic datamodule, B200-native arm and reference (CPU oracle port) arm.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--engine ffma|tc_precise|tc_fast]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

One step = gen_batch -> LSTM forward -> objective -> BPTT -> global-norm clip -> Adam on ONE batch of
`--batch` windows per GPU (default 4096 = BASELINE.json configs[1]: defaults small/mse/fast, 1xB200).
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
def oracle_steps_per_second(model_name, loss, batch_windows, steps, warmup, threads, seed=1234):
    """Times the oracle port of the reference's training step on the host cores; returns windows/s."""
    import torch
    sys.path.insert(0, str(ROOT / "oracle"))
    import lstm_oracle as O
    torch.set_num_threads(threads)
    torch.set_float32_matmul_precision("highest")
    H, L, p = MODELS[model_name]
    torch.manual_seed(seed)
    m = O.OracleLstm(3, H, L, p)
    opt = O.make_optimizer(m, LR)
    batches = [O.generate_windows(range(i * batch_windows, (i + 1) * batch_windows), seed=seed) for i in range(2)]
    for i in range(warmup):
        O.train_step(m, opt, batches[i % 2], loss, CLIP)
    t0 = time.perf_counter()
    for i in range(steps):
        O.train_step(m, opt, batches[i % 2], loss, CLIP)
    dt = time.perf_counter() - t0
    return batch_windows * steps / dt, dt / steps


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    bw = args.cpu_batch
    v, sec = oracle_steps_per_second(args.model, args.loss, bw, args.steps, args.warmup, cores)
    sample = (f"each step = one optimisation step of the oracle port on {bw} windows ({bw * S} sequences), "
              f"a 1/{max(1, args.batch // bw)} sample of the {args.batch}-window batch; torch CPU fp32 eager, {cores} threads")
    out = {"impl": "reference", "metric": "train samples/sec", "value": v, "unit": "samples/s", "n_gpus": args.gpus,
           "steps": args.steps, "warmup": args.warmup, "ms_per_step": sec * 1e3, "higher_is_better": True,
           "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
           "config": workload_config(args, 1), "cpu_baseline": {"value": v, "unit": "samples/s", "cores": cores,
                                                               "kind": "port", "sample": sample},
           "e2e": {"value": v, "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(out), flush=True)


def workload_config(args, world):
    H, L, p = MODELS[args.model]
    return {"workload": f"configs/config.yaml defaults (datamodule=synthetic model={args.model} loss={args.loss} trainer=fast "
                        f"clip {CLIP}) at datamodule.batch_size={args.batch} per GPU: {args.batch * S} sequences x {T} steps, "
                        f"H={H} L={L} dropout={p}",
            "global_batch_windows": args.batch * world, "sequences_per_step": args.batch * S * world, "lookback": T,
            "engine": args.engine, "l2": "per-step working set (activation stash, >= 25 GB at 4096 windows) exceeds the 126 MB L2",
            "parallelism": f"dp{world}"}


# --------------------------------------------------------------------------------------------- b200 arm
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
    torch.manual_seed(1234)
    cls = {"mse": B.FinancialLstmMse, "nll": B.FinancialLstmNll, "combined": B.FinancialLstmCombined}[args.loss]
    model = cls(3, H, L, p, LR)
    dm = B.FinancialLstmDataModule(lookback_window=T, target_window=N_TARGET, stride=90, batch_size=args.batch * world,
                                   seed=1234, device=dev, streaming=True, steps_per_epoch=args.steps + args.warmup,
                                   rank=rank, world_size=world)
    dm.prepare_data(); dm.setup("fit")
    tr = B.B200Trainer(max_epochs=1, gradient_clip_val=CLIP, engine=args.engine, verbose=False)
    tr._attach(model, dm)
    tr._init_distributed()
    e = model.engine
    e.ensure_capacity(args.batch)
    first = dm.n_windows

    def step(k):
        batch = dm.source.generate(first_window=first + (k * world + rank) * args.batch, n_windows=args.batch)
        return tr.training_step(model, batch, LR)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident throughput -------------------------------------------------------------
    for k in range(args.warmup):
        step(k)
    barrier()
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    l0 = e.launches + dm.source.launches
    lib.b200_timing_enable(1)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    ev0.record()
    for k in range(args.steps):
        losses = step(args.warmup + k)
    ev1.record()
    barrier()
    ms = ev0.elapsed_time(ev1)
    kt = _lib.timing_report()
    lib.b200_timing_enable(0)
    # the library's own per-name launch counts of the timed region (TimedLaunch registry); the engine's host-side counter is the fallback
    launches = int(sum(v[0] for v in kt.values())) if kt else e.launches + dm.source.launches - l0
    clk = clocks.stop() if rank == 0 else None
    t_ms = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t_ms, op=dist.ReduceOp.MAX)
    ms = float(t_ms)
    value = args.batch * world * args.steps / (ms / 1e3)
    final_loss = float(losses[2])

    # ---- end to end: reference-format batches in pinned host memory -> H2D -> step -> loss D2H ----
    e2e = None
    if not args.no_e2e:
        # two distinct batches in pinned host memory, alternated
        refs = [[t.cpu().pin_memory() for t in dm.source.generate(first_window=first + i * args.batch, n_windows=args.batch).to_reference()]
                for i in range(2)]
        h2d = sum(t.numel() * 4 for t in refs[0])
        e2e_steps = max(2, min(args.steps, 5))
        feeder = B.HostBatchFeeder(dev)

        def run_e2e(n):
            # every step's loss is read back to the host (pinned buffer + event), one step behind the enqueue front
            box, vals = B.LossMailbox(dev), []
            for b in feeder.feed(refs[i % 2] for i in range(n)):
                vals += box.post(tr.training_step(model, b, LR))
            vals += box.drain()
            assert len(vals) == n
            return vals[-1][2]

        run_e2e(2)
        barrier()
        t0 = time.perf_counter()
        run_e2e(e2e_steps)
        barrier()
        dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        e2e = {"value": args.batch * world * e2e_steps / float(dt), "unit": "samples/s", "h2d_bytes_per_step": h2d,
               "d2h_bytes_per_step": 16, "steps": e2e_steps,
               "path": "pinned host [X,y,factor,inv_psi] (reference DataLoader layout) -> HostBatchFeeder (cudaMemcpyAsync + b200_collate_reference on a copy stream, one batch ahead) -> B200Trainer.training_step -> LossMailbox (every step's [mse, nll, total, mae] copied to pinned host memory and read by the host one step behind the enqueue front)"}

    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rank != 0:
        return
    # ---- roofline of the dominant kernel (live CUDA-event durations from the timed region) -------
    hbm, tf_burst, tf_sus, src = peaks()
    dom = max(kt.items(), key=lambda kv: kv[1][1]) if kt else (None, (0, 0.0))
    seqs = args.batch * S
    F_seq = flops_per_sequence(3, H, L)
    lstm_ms = sum(v[1] for k, v in kt.items() if k.startswith(("lstm_", "wide_")))
    lstm_launch_ms = lstm_ms / args.steps if args.steps else 0.0
    achieved = seqs * F_seq / (lstm_launch_ms * 1e-3) / 1e12 if lstm_launch_ms else 0.0
    traffic = None
    tj = ROOT / "profiles" / "ncu_r01_traffic.json"
    if tj.exists():             # DRAM bytes per launch of the dominant kernel from the committed ncu --set full capture of this workload
        t = json.loads(tj.read_text())
        wl = t.get("workload", {})
        if (wl.get("model"), wl.get("engine"), wl.get("batch")) == (args.model, args.engine, args.batch):
            traffic = t["bytes_per_launch"].get(dom[0])
    roofline = {"bound": "tensor", "kernel": "LSTM forward+backward kernels (" + ", ".join(sorted(k for k in kt if k.startswith(("lstm_", "wide_")))) + ")",
                "dominant_kernel": dom[0], "achieved": achieved, "peak": tf_sus, "unit": "TFLOP/s", "frac": achieved / tf_sus,
                "traffic": traffic, "peak_source": f"{src} bf16 sustained (MEASURED_PEAKS.json)",
                "algorithmic_flops_per_step": seqs * F_seq, "lstm_ms_per_step": lstm_launch_ms,
                "kernel_ms_per_step": {k: v[1] / args.steps for k, v in sorted(kt.items())},
                "kernel_launches_per_step": {k: v[0] / args.steps for k, v in sorted(kt.items())}}
    # ---- CPU baseline: bounded sample on this box's host cores -----------------------------------
    cpu = None
    if not args.no_cpu:
        cores = os.cpu_count() or 1
        v, sec = oracle_steps_per_second(args.model, args.loss, args.cpu_batch, args.cpu_steps, 2, cores)
        cpu = {"value": v, "unit": "samples/s", "cores": cores, "kind": "port",
               "sample": f"{args.cpu_steps} optimisation steps of the oracle port at {args.cpu_batch} windows/step "
                         f"(1/{max(1, args.batch // args.cpu_batch)} of the batch), torch CPU fp32 eager, {sec:.2f} s/step"}
    out = {"metric": "train samples/sec", "value": value, "unit": "samples/s", "n_gpus": world, "steps": args.steps,
           "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
           "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": workload_config(args, world),
           "sequences_per_sec": value * S, "final_loss": final_loss, "clocks": clk, "e2e": e2e, "gpu_launches": launches,
           "roofline": roofline, "cpu_baseline": cpu}
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
    ap.add_argument("--batch", type=int, default=4096, help="windows per GPU per step")
    ap.add_argument("--cpu-batch", type=int, default=64, help="windows per step of the CPU sample")
    ap.add_argument("--cpu-steps", type=int, default=6)
    ap.add_argument("--dropout", type=float, default=None, help="override the model's inter-layer dropout (diagnostics)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    args = ap.parse_args()
    if args.dropout is not None:
        h, l, _ = MODELS[args.model]
        MODELS[args.model] = (h, l, args.dropout)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
