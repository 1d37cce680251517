"""Event timeline of CTA 0 of the tcgen05 forward kernel (diagnostics; needs a library built with B200_TRACE=1 python masters-thesis_b200/build.py --force).  python tools/trace_fwd.py [--layer 1]"""
import argparse, sys
from pathlib import Path
import numpy as np, torch
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import masters_thesis_b200 as B
from masters_thesis_b200._lib import load, check, ptr
ap = argparse.ArgumentParser(); ap.add_argument("--engine", default="tc_precise"); ap.add_argument("--batch", type=int, default=2048)
ap.add_argument("--layer", type=int, default=1); ap.add_argument("--layers", type=int, default=2); ap.add_argument("--steps", default="20,21"); ap.add_argument("--dropout", type=float, default=0.2)
a = ap.parse_args()
lib = load(); dev = "cuda:0"
torch.manual_seed(0)
# trace the chosen layer by making it the only traced launch: run a 1-layer or 2-layer model accordingly
m = B.FinancialLstmMse(3, 64, a.layers, a.dropout, 1e-4); e = m.attach(dev, a.engine)
src = B.data.DeviceBatchSource(dev); b = src.generate(first_window=0, n_windows=a.batch)
CAP = 8192
buf = torch.zeros(4 * CAP * 2, dtype=torch.int64, device=dev)
e.forward(b, True, 1); torch.cuda.synchronize()
check(lib.b200_debug_trace(ptr(buf)), "trace")
e.forward(b, True, 1); torch.cuda.synchronize()      # both layers write the trace; the later launch (layer 1) overwrites
check(lib.b200_debug_trace(None), "trace")
if a.layer == 0:
    print("note: layer 0 records are overwritten by layer 1 where both wrote; use a 1-layer model for layer 0")
rec = buf.cpu().numpy().view(np.uint64).reshape(4, CAP, 2)
names = {0: {0: "epi.wait_acc", 1: "epi.acc_ready", 6: "epi.h_ready"}, 1: {0: "ld.loaded_wait_x", 1: "ld.x_free_ok", 2: "ld.x_full"},
         2: {0: "mma.wait_x", 1: "mma.x_ok", 2: "mma.h_ok", 3: "mma.issued"}}
ev = []
for role in range(3):
    for tag, clk in rec[role]:
        if clk == 0: continue
        hi = int(tag) >> 32; step = int(tag) & 0xFFFFFFFF
        ev.append((int(clk), role, (hi >> 16) & 0xFF, (hi >> 8) & 0xFF, step))
ev.sort()
lo, hi_ = [int(x) for x in a.steps.split(",")]
sel = [x for x in ev if lo <= x[4] <= hi_]
t0 = sel[0][0]
for clk, role, e_, q, step in sel:
    print(f"{clk - t0:8d}  step {step:3d} slot{q}  {names[role].get(e_, e_)}")
ends = [x[0] for x in ev if x[1] == 0 and x[2] == 6]
print("cycles per step (median):", np.median(np.diff(ends[5:50])))
