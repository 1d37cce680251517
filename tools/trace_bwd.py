"""Event timeline of CTA 0 of the tcgen05 backward kernel (diagnostics; needs a library built with B200_TRACE=1 python masters-thesis_b200/build.py --force).  python tools/trace_bwd.py [--layer 1] [--engine tc_precise]"""
import argparse, sys
from pathlib import Path
import numpy as np, torch
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import masters_thesis_b200 as B
from masters_thesis_b200._lib import load, check, ptr

ap = argparse.ArgumentParser(); ap.add_argument("--engine", default="tc_precise"); ap.add_argument("--batch", type=int, default=512)
ap.add_argument("--layer", type=int, default=1); ap.add_argument("--steps", default="20,21"); ap.add_argument("--dropout", type=float, default=0.2)
a = ap.parse_args()
lib = load(); dev = "cuda:0"
torch.manual_seed(0)
m = B.FinancialLstmMse(3, 64, 2, a.dropout, 1e-4); e = m.attach(dev, a.engine)
src = B.data.DeviceBatchSource(dev); b = src.generate(first_window=0, n_windows=a.batch)
CAP = 8192
buf = torch.zeros(4 * CAP * 2, dtype=torch.int64, device=dev)
e.forward(b, True, 1); e.heads(b, "mse"); 
N = a.batch * 100
for layer in (1, 0):
    buf.zero_()
    check(lib.b200_debug_trace(ptr(buf) if layer == a.layer else None), "trace")
    check(lib.b200_lstm_bwd_layer(e.mode, layer, ptr(e.params), 3, 64, 2, ptr(b.x), N, 60, ptr(e.stash_h), ptr(e.stash_c), ptr(e.dhT), ptr(e.grads), e.p, 1, 1, ptr(e.lstm_ws), e.lstm_ws.numel(), None), "bwd")
    torch.cuda.synchronize()
    if layer == a.layer: rec = buf.cpu().numpy().view(np.uint64).reshape(4, CAP, 2).copy()
check(lib.b200_debug_trace(None), "trace")
names = {0: {0: "epi.wait_a", 1: "epi.a_ready", 2: "epi.rb_start", 3: "epi.rb_done", 4: "epi.wait_dafree", 5: "epi.dafree_ok", 6: "epi.da_full", 7: "epi.q_done", 8: "epi.hp1_ld_done", 9: "epi.hp1_loads_issued", 10: "epi.hp1_sts_done"},
         1: {0: "ld.begin", 1: "ld.xhfree_ok", 2: "ld.xh_full"},
         3: {4: "mmaB.wgrad_issued"},
         2: {0: "mma.wait_da", 1: "mma.da_ok", 2: "mma.dgrad_issued", 3: "mma.recomp_issued", 4: "mma.wgrad_issued", 5: "mma.next_recomp_issued"}}
ev = []
for role in range(4):
    for tag, clk in rec[role]:
        if clk == 0: continue
        hi = int(tag) >> 32; step = int(tag) & 0xFFFFFFFF
        ev.append((int(clk), role, (hi >> 16) & 0xFF, (hi >> 8) & 0xFF, step))
ev.sort()
lo, hi_ = [int(x) for x in a.steps.split(",")]
sel = [x for x in ev if lo <= x[4] <= hi_]
t0 = sel[0][0]
for clk, role, e_, q, step in sel:
    print(f"{clk - t0:8d}  step {step:3d} q{q}  {names[role][e_]}")
steps = sorted(set(x[4] for x in ev if x[1] == 0 and x[2] == 7 and x[3] == 3))
ends = {x[4]: x[0] for x in ev if x[1] == 0 and x[2] == 7 and x[3] == 3}
d = np.diff([ends[s] for s in steps][5:50])
print("cycles per step (median):", np.median(d))
