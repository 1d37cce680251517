"""clock64 stamps of CTA 0 of one w2_fwd_kernel launch (top layer, middle step; option wide_dbg bit 32, stored at the start of the unused
stash_h): where the epilogue, the MMA issuer and the loader spend a tile's time.   python tools/wide2_trace.py [windows] [layers]"""
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parents[1]
sys.path[:0] = [str(ROOT)]
import masters_thesis_b200 as B  # noqa: E402
from masters_thesis_b200.data import DeviceBatchSource  # noqa: E402


def main():
    windows = int(sys.argv[1]) if len(sys.argv) > 1 else 512
    layers = int(sys.argv[2]) if len(sys.argv) > 2 else 2
    dev = "cuda:0"
    torch.manual_seed(1)
    e = B.LstmEngine(3, 256, layers, 0.3, dev, "tc_fast")
    e.load_params(torch.randn(e.P) * 0.05)
    batch = DeviceBatchSource(dev, seed=5).generate(first_window=0, n_windows=windows)
    e.forward(batch, training=True, dropout_seed=3)           # warm-up, allocates
    assert e.lib.b200_set_option(b"wide_dbg", 32) == 0
    e.stash_h[: 3 * 256 * 4 * 2].zero_()
    e.forward(batch, training=True, dropout_seed=3)
    torch.cuda.synchronize()
    e.lib.b200_set_option(b"wide_dbg", 0)
    tr = e.stash_h[: 3 * 256 * 4 * 2].view(torch.int64).cpu().reshape(3, 256, 4)
    n = int((tr[1, :, 3] > 0).sum())
    t0 = int(tr[1, 0, 0])
    print(f"tiles of CTA 0: {n}; clocks relative to the MMA issuer's first stamp")
    print("tile | MMA: wait acc_free, wait 1st chunk, issue+waits, period | epilogue: start-wait, wait acc, work | loader: tile start, 1st issued, all issued")
    prev = None
    for i in range(n):
        m = [int(x) - t0 for x in tr[1, i]]
        ep = [int(x) - t0 for x in tr[0, i, :3]]
        ld = [int(x) - t0 for x in tr[2, i, :3]]
        period = m[3] - prev if prev is not None else 0
        prev = m[3]
        print(f"{i:3d} | {m[1]-m[0]:6d} {m[2]-m[1]:6d} {m[3]-m[2]:6d} {period:6d} | {ep[0]:7d} {ep[1]-ep[0]:6d} {ep[2]-ep[1]:6d} | {ld[0]:7d} {ld[1]-ld[0]:5d} {ld[2]-ld[0]:6d}")


if __name__ == "__main__":
    main()
