"""Bring-up check of the bf16 wide engine (csrc/lstm_wide2.cu): hidden states of every layer, loss and per-tensor gradients of
B200_MODE_TC_FAST against B200_MODE_TC_PRECISE (the fp32-storage engine, itself gated against the oracle) on the golden batches.

    python tools/wide2_check.py [H] [layers] [dropout] [windows] [option=value ...]
"""
import json
import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parents[1]
sys.path[:0] = [str(ROOT), str(ROOT / "tests"), str(ROOT / "oracle")]
import lstm_oracle as O  # noqa: E402
import masters_thesis_b200 as B  # noqa: E402
from _gpu_util import DEV, golden_batches, make_engine, ref_batch, rel_err  # noqa: E402


def main():
    H = int(sys.argv[1]) if len(sys.argv) > 1 else 256
    layers = int(sys.argv[2]) if len(sys.argv) > 2 else 2
    p = float(sys.argv[3]) if len(sys.argv) > 3 else 0.0
    nwin = int(sys.argv[4]) if len(sys.argv) > 4 else 2
    for o in sys.argv[5:]:                      # library options name=value (b200_set_option)
        k, v = o.split("=")
        assert B._lib.load().b200_set_option(k.encode(), int(v)) == 0, o
    data = golden_batches(ROOT / "tests" / "golden")
    torch.manual_seed(4)
    m = O.OracleLstm(3, H, layers, 0.0)
    flat = O.flat_params(m).clone()
    batch = B.DeviceBatch.from_reference(ref_batch(data, 0, nwin), DEV)
    out = {}
    res = {}
    for mode in ("tc_precise", "tc_fast"):
        e = make_engine(flat, layers, p, mode, 3, H)
        e.forward(batch, training=True, dropout_seed=77)
        hs = [e.hidden_states(l, nwin).clone() for l in range(layers)]
        losses = e.heads(batch, "mse", 100.0, None, 1.0, True).clone()
        e.backward(batch, True, dropout_seed=77)
        torch.cuda.synchronize()
        res[mode] = (hs, losses.cpu().numpy(), e.grads[: e.P].cpu().numpy().copy())
    for l in range(layers):
        a, b = res["tc_fast"][0][l], res["tc_precise"][0][l]
        out[f"h{l}_maxabs"] = float((a - b).abs().max())
        out[f"h{l}_t0_maxabs"] = float((a[0] - b[0]).abs().max())
    out["loss_fast"], out["loss_precise"] = float(res["tc_fast"][1][2]), float(res["tc_precise"][1][2])
    got, want = res["tc_fast"][2], res["tc_precise"][2]
    out["grad_rel"] = rel_err(got, want)
    out["grad_cos"] = float(np.dot(got, want) / (np.linalg.norm(got) * np.linalg.norm(want) + 1e-30))
    off = 0
    for name, p_ in m.named_parameters():
        n = p_.numel()
        out["g_" + name] = round(rel_err(got[off:off + n], want[off:off + n]), 5)
        off += n
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
