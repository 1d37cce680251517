import sys, numpy as np, torch
sys.path.insert(0, "."); sys.path.insert(0, "tests"); sys.path.insert(0, "oracle")
import lstm_oracle as O, masters_thesis_b200 as B
from _gpu_util import *
from pathlib import Path
gd = Path("tests/golden")
data = golden_batches(gd)
for kind in ("mse", "nll", "combined"):
    g = np.load(gd / f"golden_{kind}.npz")
    e = make_engine(torch.from_numpy(g["init"]), 2, 0.0, "tc_precise")
    batch = B.DeviceBatch.from_reference(ref_batch(data, 0, 2), DEV)
    e.forward(batch, training=True); e.heads(batch, kind, 100.0, None, 1.0, True); e.backward(batch, True)
    got = e.grads[: e.P].cpu().numpy(); off = 0
    m = O.OracleLstm(3, 64, 2, 0.0)
    out = {}
    for name, p_ in m.named_parameters():
        n = p_.numel(); out[name] = rel_err(got[off:off + n], g["grads0"][off:off + n]); off += n
    print(kind, {k: f"{v:.1e}" for k, v in out.items() if "l1" in k or "l0" in k})
# larger batch, 4 layers, dropout: vs oracle with shared masks
torch.manual_seed(11)
m = O.OracleLstm(3, 64, 4, 0.0); e = make_engine(O.flat_params(m), 4, 0.3, "tc_precise")
seed, N = 0xABC, 600
X, y, f, ip = ref_batch(data, 0, 6)
batch = B.DeviceBatch.from_reference([X, y, f, ip], DEV)
e.forward(batch, training=True, dropout_seed=seed); e.heads(batch, "combined", 100.0, None, 1.0, True); e.backward(batch, True, dropout_seed=seed)
masks = [torch.from_numpy(O.dropout_mask(seed, l + 1, N, 60, 64, 0.3)).permute(1, 0, 2) for l in range(3)]
em = O.ExplicitMaskLstm(m)
ls = O.batch_losses(m, [X, y, f, ip], "combined", 100.0, forward=lambda x: em(x, masks)); ls["total"].backward()
got = e.grads[: e.P].cpu().numpy(); want = O.flat_grads(m).numpy(); off = 0
for name, p_ in m.named_parameters():
    n = p_.numel(); print(name, f"{rel_err(got[off:off+n], want[off:off+n]):.1e}", end="; "); off += n
print()
