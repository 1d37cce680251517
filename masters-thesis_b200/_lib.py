"""ctypes binding of libb200lstm.so (include/b200_lstm.h).  No CPU fallback: if the library is
missing or a call fails this raises -- the product path never routes around the CUDA kernels."""
from __future__ import annotations

import ctypes as C
from pathlib import Path

PKG = Path(__file__).resolve().parent
LIB_PATH = PKG / "libb200lstm.so"

MODE_FFMA, MODE_TC_PRECISE, MODE_TC_FAST = 0, 1, 2
MODES = {"ffma": MODE_FFMA, "tc_precise": MODE_TC_PRECISE, "tc_fast": MODE_TC_FAST}
LOSS_KINDS = {"mse": 0, "nll": 1, "combined": 2}

_p, _i, _l, _f, _u64, _sz = C.c_void_p, C.c_int, C.c_int64, C.c_float, C.c_uint64, C.c_size_t

# name -> (restype, argtypes); kept in the order of include/b200_lstm.h
PROTOTYPES = {
    "b200_version": (_i, []),
    "b200_last_error": (C.c_char_p, []),
    "b200_build_hash": (C.c_char_p, []),
    "b200_mode_available": (_i, [_i]),
    "b200_timing_enable": (_i, [_i]),
    "b200_timing_report": (_i, [C.c_char_p, _sz]),
    "b200_set_option": (_i, [C.c_char_p, _i]),
    "b200_param_count": (_l, [_i, _i, _i]),
    "b200_param_offset": (_l, [_i, _i, _i, _i]),
    "b200_gen_batch": (_i, [_u64, _p, _l, _l, _i, _i, _i, _l, _i, _p, _p, _p, _p, _p, _p]),
    "b200_series_batch": (_i, [_p, _p, _l, _p, _l, _l, _i, _i, _i, _l, _i, _i, _p, _p, _p, _p, _p, _p, _p, _p]),
    "b200_gen_series": (_i, [_u64, _l, _l, _i, _p, _p, _p]),
    "b200_collate_reference": (_i, [_p, _p, _l, _i, _i, _i, _i, _p, _p, _p, _p]),
    "b200_gen_coefficients": (_i, [_u64, _i, _p, _p, _p]),
    "b200_philox4x32": (_i, [_u64, _p, _l, _p, _p]),
    "b200_dropout_mask": (_i, [_u64, _i, _l, _i, _i, _f, _p, _p]),
    "b200_lstm_workspace_bytes": (_sz, [_i, _l, _i, _i, _i, _i]),
    "b200_lstm_fwd": (_i, [_i, _p, _i, _i, _i, _p, _l, _i, _p, _p, _f, _u64, _i, _p, _sz, _p]),
    "b200_lstm_bwd_layer": (_i, [_i, _i, _p, _i, _i, _i, _p, _l, _i, _p, _p, _p, _p, _f, _u64, _i, _p, _sz, _p]),
    "b200_lstm_bwd": (_i, [_i, _p, _i, _i, _i, _p, _l, _i, _p, _p, _p, _p, _f, _u64, _i, _p, _sz, _p]),
    "b200_lstm_hidden_states": (_i, [_i, _i, _i, _i, _i, _l, _i, _p, _p, _sz, _p, _p]),
    "b200_tc_selftest": (_i, [_p, _p, _i, _i, _i, _p, _p]),
    "b200_tc_mma_bench": (_i, [_i, _i, _i, _i, _p, _p]),
    "b200_wide_gemm": (_i, [_i, _i, _p, _p, _p, _l, _i, _l, _i, _p, _sz, _p]),
    "b200_debug_trace": (_i, [_p]),
    "b200_heads_workspace_bytes": (_sz, [_l, _i, _i, _i]),
    "b200_heads_loss_fwd_bwd": (_i, [_i, _f, _p, _i, _p, _p, _p, _p, _p, _l, _i, _i, _l, _f, _p, _p, _p, _p, _p, _p, _sz, _p]),
    "b200_heads_mse_fwd_bwd": (_i, [_p, _i, _p, _p, _p, _l, _i, _i, _f, _p, _p, _p, _p, _sz, _p]),
    "b200_heads_nll_fwd_bwd": (_i, [_p, _i, _p, _p, _p, _p, _l, _i, _i, _l, _f, _p, _p, _p, _p, _sz, _p]),
    "b200_sqnorm_workspace_bytes": (_sz, [_l]),
    "b200_grad_sqnorm": (_i, [_p, _l, _p, _p, _sz, _p]),
    "b200_grad_sqnorm_static": (_i, [_p, _l, _p, _p]),
    "b200_clip_adam_step": (_i, [_p, _p, _p, _p, _l, _f, _f, _f, _f, _f, _l, _f, _p, _p]),
    "b200_grad_accumulate": (_i, [_p, _p, _l, _f, _p]),
    "b200_adam_step": (_i, [_p, _p, _p, _p, _l, _f, _f, _f, _f, _f, _l, _f, _p, _i, _p]),
    "b200_dp_nccl_version": (_i, []),
    "b200_dp_unique_id": (_i, [_p]),
    "b200_dp_init": (_i, [_p, _i, _i, C.POINTER(_p)]),
    "b200_dp_allreduce": (_i, [_p, _p, _l, _p]),
    "b200_dp_destroy": (_i, [_p]),
}


class B200Error(RuntimeError):
    pass


_lib = None


def load(path: Path | None = None):
    """Load the shared library once and attach prototypes.  Raises if it has not been built."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    path = Path(path) if path else LIB_PATH
    if not path.exists():
        raise B200Error(f"{path} not found: build it with `python masters-thesis_b200/build.py` "
                        "(nvcc, sm_100a).  There is no CPU fallback.")
    lib = C.CDLL(str(path))
    for name, (res, args) in PROTOTYPES.items():
        fn = getattr(lib, name)       # AttributeError if the symbol is not exported
        fn.restype, fn.argtypes = res, args
    # a prebuilt binary that does not match the sources next to it must not pass silently
    from . import build as _build
    if _build.CSRC.exists():
        want, got = _build.source_hash(_build._extra_flags()), lib.b200_build_hash().decode()
        if got != want:
            raise B200Error(f"{path} was built from other sources (library {got}, tree {want}): "
                            "run `python masters-thesis_b200/build.py`")
    _lib = lib
    return lib


def check(rc: int, what: str):
    if rc != 0:
        msg = load().b200_last_error().decode(errors="replace")
        raise B200Error(f"{what} failed with code {rc}: {msg}")


def ptr(t):
    """Device pointer of a torch tensor (None -> NULL)."""
    return None if t is None else C.c_void_p(t.data_ptr())


def timing_report():
    """{kernel name: (launches, total_ms)} accumulated since b200_timing_enable(1)."""
    buf = C.create_string_buffer(1 << 14)
    check(load().b200_timing_report(buf, len(buf)), "b200_timing_report")
    out = {}
    for line in buf.value.decode().splitlines():
        name, n, ms = line.rsplit(" ", 2)
        out[name] = (int(n), float(ms))
    return out
