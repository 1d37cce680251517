"""One forward + backward of the widened model (bf16 operand-image engine) on a device-generated batch, for ncu:

    ncu --set full --clock-control none -k regex:w2_fwd_kernel --launch-skip 90 --launch-count 1 -o gpurun_out/w2_fwd \
        python tools/wide2_profile.py 512 2
"""
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parents[1]
sys.path[:0] = [str(ROOT)]
import masters_thesis_b200 as B  # noqa: E402


def main():
    windows = int(sys.argv[1]) if len(sys.argv) > 1 else 512
    layers = int(sys.argv[2]) if len(sys.argv) > 2 else 2
    reps = int(sys.argv[3]) if len(sys.argv) > 3 else 1
    dev = "cuda:0"
    torch.manual_seed(1)
    e = B.LstmEngine(3, 256, layers, 0.3, dev, "tc_fast")
    e.load_params(torch.randn(e.P) * 0.05)
    from masters_thesis_b200.data import DeviceBatchSource
    src = DeviceBatchSource(dev, seed=5)
    batch = src.generate(first_window=0, n_windows=windows)
    for _ in range(reps):
        e.forward(batch, training=True, dropout_seed=3)
        e.heads(batch, "mse", 100.0, None, 1.0, True)
        e.backward(batch, True, dropout_seed=3)
        e.optimizer_step(1e-4, 5.0)          # global-norm clip + Adam over the flat vector (3.95 M parameters at 8 layers)
    torch.cuda.synchronize()
    print("ok", float(e.losses[2]))


if __name__ == "__main__":
    main()
