"""Diagnostic (GPU box): per-tensor gradient error of one full-size step (4096 windows, small model, mse, dropout p) of the
tc_precise and ffma engines against (a) the fp32 CPU oracle accumulated over 256-window chunks (fp32 and fp64 accumulation
across chunks) and (b) an fp64 evaluation of the same network with torch on the GPU (truth).  Prints one JSON line.
    python tools/fullsize_grad_diag.py [windows] [dropout]"""
import json, os, sys
from pathlib import Path
import numpy as np, torch
ROOT = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(ROOT), str(ROOT / "oracle"), str(ROOT / "tests")]
import lstm_oracle as O
import masters_thesis_b200 as B
from _budget import device_masks, S

from masters_thesis_b200._lib import load
if len(sys.argv) > 3:          # library option for the device engines, e.g. tc_wgrad=2
    k, _, v = sys.argv[3].partition("="); load().b200_set_option(k.encode(), int(v))
BW = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
P = float(sys.argv[2]) if len(sys.argv) > 2 else 0.2
dev, seed = "cuda:0", 1234
torch.set_num_threads(os.cpu_count()); torch.set_float32_matmul_precision("highest")
torch.manual_seed(7)
model = B.FinancialLstmMse(3, 64, 2, P, 1e-4)
flat = model.flat_parameters().clone()
src = B.data.DeviceBatchSource(dev, seed)
N = BW * S
dseed = 4242
grads = {}
for eng in ("tc_precise", "ffma"):
    e = B.LstmEngine(3, 64, 2, P, dev, eng); e.load_params(flat)
    batch = src.generate(first_window=20000, n_windows=BW)
    e.forward_backward(batch, "mse", dropout_seed=dseed)
    torch.cuda.synchronize()
    grads[eng] = e.grads[: e.P].double().cpu().numpy()
    del e; torch.cuda.empty_cache()
masks = device_masks(dseed, 2, N, P, dev) if P > 0 else None

def build(dtype, device):
    om = O.OracleLstm(3, 64, 2, 0.0).to(dtype)
    off = 0
    with torch.no_grad():
        for p_ in om.parameters():
            p_.copy_(flat[off:off + p_.numel()].reshape(p_.shape).to(dtype)); off += p_.numel()
    return om.to(device)

def run(dtype, device, chunk, acc64):
    om = build(dtype, device); em = O.ExplicitMaskLstm(om)
    acc = np.zeros(flat.numel(), np.float64); loss = 0.0
    om.zero_grad()
    for lo in range(0, BW, chunk):
        nb = min(chunk, BW - lo)
        ref = [t.to(device=device, dtype=dtype) for t in src.generate(first_window=20000 + lo, n_windows=nb).to_reference()]
        fwd = None
        if masks:
            mk = [m[:, lo * S:(lo + nb) * S].permute(1, 0, 2).to(device=device, dtype=dtype) for m in masks]
            fwd = lambda x, _mk=mk: em(x, _mk)
        if acc64: om.zero_grad()
        ls = O.batch_losses(om, ref, "mse", forward=fwd)
        (ls["total"] * (nb / BW)).backward()
        loss += float(ls["total"].detach()) * nb / BW
        if acc64: acc += O.flat_grads(om).double().cpu().numpy()
    return (acc if acc64 else O.flat_grads(om).double().cpu().numpy()), loss

quick = len(sys.argv) > 3           # with an option: only the fp64 truth (the CPU oracle legs take a minute each at full size)
refs = {"gpu64": run(torch.float64, dev, min(512, BW), True)[0]}
if not quick:
    refs.update({"cpu32_acc32": run(torch.float32, "cpu", 256, False)[0], "cpu32_acc64": run(torch.float32, "cpu", 256, True)[0]})
names, shapes = model.names, B.model._param_shapes(3, 64, 2)
def per_tensor(a, b):
    out, off = {}, 0
    for n_, s_ in zip(names, shapes):
        k = int(np.prod(s_)); out[n_] = float(np.abs(a[off:off+k] - b[off:off+k]).max() / (np.abs(b[off:off+k]).max() + 1e-30)); off += k
    out["whole"] = float(np.abs(a - b).max() / np.abs(b).max())
    return out
res = {"windows": BW, "dropout": P}
for a in ("tc_precise", "ffma", "cpu32_acc32", "cpu32_acc64"):
    if a in grads or a in refs:
        res[f"{a}_vs_gpu64"] = per_tensor(grads[a] if a in grads else refs[a], refs["gpu64"])
if not quick:
    res["tc_precise_vs_cpu32_acc32"] = per_tensor(grads["tc_precise"], refs["cpu32_acc32"])
    res["tc_precise_vs_cpu32_acc64"] = per_tensor(grads["tc_precise"], refs["cpu32_acc64"])
res["option"] = sys.argv[3] if quick else None
print(json.dumps(res))
