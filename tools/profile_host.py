"""Host-side profile of the B = 1 training loop (reference default config): where do the 2.4 ms per step go?"""
import cProfile, pstats, sys, time
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import masters_thesis_b200 as B
torch.manual_seed(0)
model = B.FinancialLstmMse(3, 64, 2, 0.2, 1e-4)
dm = B.FinancialLstmDataModule(lookback_window=60, target_window=30, stride=90, batch_size=int(sys.argv[1]) if len(sys.argv) > 1 else 1,
                               seed=1234, n_samples=90 * 2000)
tr = B.B200Trainer(max_epochs=1, gradient_clip_val=5.0, engine="auto", verbose=False, limit_val_windows=4)
tr.fit(model, dm)          # warm-up epoch (1400 steps)
torch.cuda.synchronize()
pr = cProfile.Profile()
t0 = time.time(); pr.enable()
tr.fit(model, dm)
torch.cuda.synchronize(); pr.disable()
dt = time.time() - t0
print(f"{tr.global_step // 2} steps/epoch, {1e3 * dt / (tr.global_step // 2):.3f} ms/step")
pstats.Stats(pr).sort_stats("cumulative").print_stats(22)
