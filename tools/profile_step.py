"""Runs N training steps of the bench workload (for ncu captures).  python tools/profile_step.py [--batch 4096] [--steps 1] [--engine tc_precise]"""
import argparse, sys
from pathlib import Path
import torch
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import masters_thesis_b200 as B

ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, default=4096); ap.add_argument("--steps", type=int, default=1)
ap.add_argument("--engine", default="tc_precise"); ap.add_argument("--dropout", type=float, default=0.2)
ap.add_argument("--hidden", type=int, default=64); ap.add_argument("--layers", type=int, default=2)
a = ap.parse_args()
dev = "cuda:0"
torch.manual_seed(0)
m = B.FinancialLstmMse(3, a.hidden, a.layers, a.dropout, 1e-4)
e = m.attach(dev, a.engine)
src = B.data.DeviceBatchSource(dev)
for k in range(a.steps):
    b = src.generate(first_window=k * a.batch, n_windows=a.batch)
    out = e.train_step(b, "mse", 1e-4, 1.0, dropout_seed=k + 1)
torch.cuda.synchronize()
print("ok", out.tolist())
