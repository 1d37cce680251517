"""Helpers shared by the -m gpu tests: everything goes through the C-ABI (ctypes) of libb200lstm.so."""
import numpy as np
import torch

import lstm_oracle as O
import masters_thesis_b200 as B

DEV = "cuda:0"


def golden_batches(golden_dir):
    d = np.load(golden_dir / "golden_data.npz")
    return {k: torch.from_numpy(d[k]) for k in d.files}


def ref_batch(data, lo, hi):
    return [data["X"][lo:hi], data["y"][lo:hi], data["factor"][lo:hi], data["inv_psi"][lo:hi]]


def make_engine(flat_params, layers=2, dropout=0.0, mode="ffma", F=3, H=64):
    e = B.LstmEngine(F, H, layers, dropout, DEV, mode)
    e.load_params(flat_params)
    return e


def oracle_model(flat_params, layers=2, dropout=0.0, F=3, H=64):
    m = O.OracleLstm(F, H, layers, dropout)
    off = 0
    with torch.no_grad():
        for p in m.parameters():
            n = p.numel()
            p.copy_(flat_params[off:off + n].reshape(p.shape))
            off += n
    return m


def rel_err(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return float(np.abs(a - b).max() / (np.abs(b).max() + 1e-30))
