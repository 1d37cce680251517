"""-m gpu, collected last: the allocation-free entry point b200_grad_sqnorm_static.

It exists for callers that cannot give the library a workspace -- a binding that only holds the
gradient pointer, or a loop whose framework has already claimed the device memory so that a new
cudaMalloc could fail or stall the GPU.  Its scratch is an 8 KB array inside the library, shared by all
calls on a device, so the test calls it twice in a row on different inputs (the second call must not
see anything left over from the first) and compares each result with b200_grad_sqnorm, through the
C-ABI.  Runs correctly alone in a fresh process and after the rest of the suite.
"""
import pytest
import torch

import masters_thesis_b200 as B
from masters_thesis_b200._lib import check, load, ptr

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
STATIC_MAX = 4194304


def test_static_scratch_matches_workspace_version_twice_in_a_row():
    lib = load()
    g = torch.Generator().manual_seed(31)
    ws = torch.empty(lib.b200_sqnorm_workspace_bytes(STATIC_MAX), dtype=torch.uint8, device=DEV)
    got, want = torch.zeros(1, device=DEV), torch.zeros(1, device=DEV)
    # first call: a large vector that fills many scratch slots; second: a short one with a ragged tail
    for count, scale in [(STATIC_MAX, 3.0), (51074, 0.01), (5, 1.0)]:
        x = (torch.randn(count, generator=g) * scale).to(DEV)
        pad = torch.zeros((count + 3) // 4 * 4, device=DEV)
        pad[:count] = x
        check(lib.b200_grad_sqnorm_static(ptr(pad), count, ptr(got), None), "b200_grad_sqnorm_static")
        check(lib.b200_grad_sqnorm(ptr(pad), count, ptr(want), ptr(ws), ws.numel(), None), "b200_grad_sqnorm")
        torch.cuda.synchronize()
        assert float(got) == pytest.approx(float(want), rel=1e-5)          # tolerance of tests/test_gpu_kernels.py
        assert float(got) == pytest.approx(float((x.double() ** 2).sum()), rel=1e-5)


def test_static_scratch_on_a_side_stream():
    """The same comparison with both calls enqueued on a non-default stream (the caller's stream is honoured)."""
    lib = load()
    side = torch.cuda.Stream(device=DEV)
    x = torch.randn(117 * 1024, generator=torch.Generator().manual_seed(5)).to(DEV)
    ws = torch.empty(lib.b200_sqnorm_workspace_bytes(x.numel()), dtype=torch.uint8, device=DEV)
    got, want = torch.zeros(1, device=DEV), torch.zeros(1, device=DEV)
    side.wait_stream(torch.cuda.current_stream())
    for _ in range(2):
        check(lib.b200_grad_sqnorm_static(ptr(x), x.numel(), ptr(got), side.cuda_stream), "b200_grad_sqnorm_static")
        check(lib.b200_grad_sqnorm(ptr(x), x.numel(), ptr(want), ptr(ws), ws.numel(), side.cuda_stream), "b200_grad_sqnorm")
    side.synchronize()
    assert float(got) == pytest.approx(float(want), rel=1e-5)
    assert float(got) == pytest.approx(float((x.double() ** 2).sum()), rel=1e-5)


def test_static_scratch_rejects_oversize_and_bad_arguments():
    lib = load()
    x = torch.zeros(8, device=DEV)
    out = torch.zeros(1, device=DEV)
    rc = lib.b200_grad_sqnorm_static(ptr(x), STATIC_MAX + 1, ptr(out), None)
    assert rc == -1 and b"exceed the built-in scratch" in lib.b200_last_error()
    assert lib.b200_grad_sqnorm_static(None, 8, ptr(out), None) == -1
    assert lib.b200_grad_sqnorm_static(ptr(x), 0, ptr(out), None) == -1
