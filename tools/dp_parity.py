"""Data-parallel parity: K optimisation steps of the mse objective on the same GLOBAL batches with
W ranks (torchrun, NCCL, bucketed all-reduce overlapped with backward) and with one rank; final parameters must agree.

    torchrun --nproc-per-node 2 tools/dp_parity.py --out gpurun_out/dp_w2.pt
    python tools/dp_parity.py --out gpurun_out/dp_w1.pt
    python tools/dp_parity.py --compare gpurun_out/dp_w1.pt gpurun_out/dp_w2.pt
"""
import argparse, os, sys
from pathlib import Path
import torch
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

ap = argparse.ArgumentParser()
ap.add_argument("--steps", type=int, default=6); ap.add_argument("--global-batch", type=int, default=512)
ap.add_argument("--engine", default="tc_precise"); ap.add_argument("--out"); ap.add_argument("--compare", nargs=2)
a = ap.parse_args()
if a.compare:
    x, y = torch.load(a.compare[0]), torch.load(a.compare[1])
    dp = (x["params"] - y["params"]).abs().max().item()
    dl = max(abs(p - q) / abs(p) for p, q in zip(x["epoch_mse"], y["epoch_mse"]))
    print({"world": (x["world"], y["world"]), "max_abs_param_diff": dp, "max_rel_epoch_loss_diff": dl, "steps": x["steps"],
           "global_batch": x["global_batch"]})
    assert dp < 5e-6 and dl < 1e-5
    sys.exit(0)

import masters_thesis_b200 as B
world, rank, local = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0"))
torch.manual_seed(1234)
model = B.FinancialLstmMse(3, 64, 2, 0.0, 1e-4)       # dropout 0: masks are per rank-local row, not per global row
dm = B.FinancialLstmDataModule(lookback_window=60, target_window=30, stride=90, batch_size=a.global_batch // world, seed=1234,      # batch_size is per rank (Lightning DDP semantics)
                              
                               device=f"cuda:{local}", streaming=True, steps_per_epoch=2)
tr = B.B200Trainer(max_epochs=a.steps // 2, gradient_clip_val=5.0, engine=a.engine, verbose=False, limit_val_windows=8)
tr.fit(model, dm)
if rank == 0:
    torch.save({"params": model.flat_parameters(), "epoch_mse": [h["loss/mse/train"] for h in tr.history], "world": world,
                "steps": tr.global_step, "global_batch": a.global_batch}, a.out)
    print("world", world, "steps", tr.global_step, [h["loss/mse/train"] for h in tr.history])
if world > 1:
    import torch.distributed as dist
    dist.barrier(); dist.destroy_process_group()
