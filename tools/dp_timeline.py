"""Where the gradient all-reduces sit in time relative to the backward kernels (SURVEY.md section 8d "evidence", in place of an
nsys timeline: nsys is not installed).  Run under torchrun on N >= 2 GPUs:

    python -m torch.distributed.run --nproc-per-node N tools/dp_timeline.py [--batch 4096] [--steps 6] --out timeline.json

Every rank runs data-parallel training steps (small model, mse, tc_precise) and brackets, with CUDA events, each backward
layer on the compute stream and each b200_dp_allreduce on the communication stream; all times are milliseconds after the
step's first kernel.  An all-reduce "overlaps" when it ends before the layer-0 backward kernel does."""
import argparse, json, os, sys
from pathlib import Path
import torch
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import masters_thesis_b200 as B

ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, default=4096); ap.add_argument("--steps", type=int, default=6)
ap.add_argument("--bucket-mb", type=float, default=0.05); ap.add_argument("--out", default=None)
a = ap.parse_args()
world, rank, local = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0"))
dev = torch.device(f"cuda:{local}")
torch.cuda.set_device(dev)
torch.manual_seed(1234)
model = B.FinancialLstmMse(3, 64, 2, 0.2, 1e-4)
dm = B.FinancialLstmDataModule(lookback_window=60, target_window=30, stride=90, batch_size=a.batch, seed=1234, device=dev,
                               streaming=True, steps_per_epoch=a.steps, rank=rank, world_size=world)
dm.prepare_data(); dm.setup("fit")
tr = B.B200Trainer(max_epochs=1, gradient_clip_val=5.0, engine="tc_precise", verbose=False, bucket_mb=a.bucket_mb)
tr._attach(model, dm); tr._init_distributed()
e, sync = model.engine, tr._sync
marks = []                                   # (name, start event, end event) of the current step

def ev(stream):
    x = torch.cuda.Event(enable_timing=True); x.record(stream); return x

orig_bwd, orig_flush = e.backward_layer, sync._flush
def bwd(batch, layer, *args, **kw):
    cur = torch.cuda.current_stream(dev)
    s = ev(cur); orig_bwd(batch, layer, *args, **kw); marks.append((f"backward layer {layer} (3 launches)", s, ev(cur)))
def flush():
    if sync._hi is None or sync._hi == sync._lo: return
    n = (sync._hi - sync._lo) * 4
    # the communication stream first waits for the compute-stream event inside _flush; events around the call on that stream
    s1 = torch.cuda.Event(enable_timing=True)
    cur = torch.cuda.current_stream(dev)
    ready = ev(cur)                           # the slice's last kernel is queued: the bucket is ready here
    orig_flush()
    s1.record(sync.comm_stream)
    marks.append((f"all-reduce {n} B ready", ready, ready)); marks.append((f"all-reduce {n} B done", s1, s1))
e.backward_layer, sync._flush = bwd, flush
steps = []
for k, batch in enumerate(dm.train_dataloader()):
    marks.clear()
    t0 = ev(torch.cuda.current_stream(dev))
    tr.training_step(model, batch, 1e-4)
    t1 = ev(torch.cuda.current_stream(dev))
    torch.cuda.synchronize()
    steps.append({"step_ms": t0.elapsed_time(t1), "events": [{"what": n, "start_ms": round(t0.elapsed_time(s), 3), "end_ms": round(t0.elapsed_time(f), 3)} for n, s, f in marks]})
import torch.distributed as dist
out = [None] * world
dist.all_gather_object(out, steps[-1])
if rank == 0:
    last = out[0]
    l0_end = max(x["end_ms"] for x in last["events"] if x["what"].startswith("backward layer 0"))
    ar = [x for x in last["events"] if x["what"].endswith("done")]
    text = json.dumps({"tool": "tools/dp_timeline.py", "world": world, "batch_windows_per_gpu": a.batch, "bucket_mb": a.bucket_mb,
                      "what": "last of %d steps, every rank; times in ms after the step's first kernel" % a.steps,
                      "rank0_layer0_backward_end_ms": l0_end,
                      "rank0_allreduce_done_ms": [x["end_ms"] for x in ar],
                      "rank0_exposed_after_backward_ms": round(max(0.0, max(x["end_ms"] for x in ar) - l0_end), 3),
                      "ranks": out})
    if a.out:
        Path(a.out).write_text(text)
    else:
        print(text)
dist.barrier(); dist.destroy_process_group()
