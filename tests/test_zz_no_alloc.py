"""hpc_colsum_static: the allocation-free variant of hpc_colsum (scratch in a static device array inside the library)
-- no counterpart in the reference.  Called through the C ABI and checked
against the existing entry point hpc_colsum (caller-provided scratch) on two calls in a row.

Both entries run the same two kernels, so the tolerance is the tightest the suite uses: bit equality.
This file sorts last on purpose (pytest collects alphabetically)."""
import ctypes

import pytest
import torch

from image_compression_b200 import _lib
from image_compression_b200._lib import check, lib, ptr, stream

pytestmark = pytest.mark.gpu
DEV = "cuda"
MAX_C = 512          # HPC_COLSUM_STATIC_MAX_C (include/hpcodec.h)


def _colsum_ref(x, scale):
    """the existing entry point, scratch from the caller"""
    P, C = x.shape
    out = torch.empty(C, dtype=torch.float32, device=x.device)
    nb = lib().hpc_colsum_scratch_bytes(P, C)
    sc = torch.empty(nb, dtype=torch.uint8, device=x.device)
    check(lib().hpc_colsum(ptr(x), _lib.dtype_code(x), P, C, scale, ptr(out), ptr(sc), nb, stream()))
    return out


def _colsum_static(x, scale, out):
    P, C = x.shape
    return lib().hpc_colsum_static(ptr(x), _lib.dtype_code(x), P, C, scale, ptr(out), stream())


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
def test_static_scratch_colsum_matches_hpc_colsum_on_two_calls_in_a_row(dtype):
    g = torch.Generator(device=DEV).manual_seed(41)
    # call 1 fills all 592 partial rows at the widest supported C; call 2 is smaller in both directions (fewer partial
    # rows, fewer columns), so anything left over from call 1 sits right next to what call 2 reads; call 3 = call 1's
    # shape again with other data
    shapes = [(70001, MAX_C), (37, 192), (70001, MAX_C)]
    xs = [(torch.randn(P, C, device=DEV, generator=g) * 3).to(dtype) for P, C in shapes]
    outs = [torch.full((C,), float("nan"), dtype=torch.float32, device=DEV) for _, C in shapes]
    scales = (1.0, -0.5, 2.0)
    refs = [_colsum_ref(x, s) for x, s in zip(xs, scales)]
    # set-up call: the library resolves the array's device address once per device (cudaGetSymbolAddress); done here so
    # that the measured region below holds steady-state calls only
    check(_colsum_static(xs[1], 1.0, torch.empty(192, dtype=torch.float32, device=DEV)))
    torch.cuda.synchronize()
    used0 = torch.cuda.memory_allocated()
    free0 = torch.cuda.mem_get_info()[0]
    for x, s, o in zip(xs, scales, outs):                    # back to back on one stream, no synchronisation between
        check(_colsum_static(x, s, o))
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == used0            # nothing allocated through torch ...
    assert torch.cuda.mem_get_info()[0] == free0             # ... nor behind its back (driver-level free memory unchanged)
    for o, r, x, s in zip(outs, refs, xs, scales):
        assert torch.equal(o, r)                             # same kernels, same order of summation: bit equal
        want = x.double().sum(0) * s                         # and both agree with a plain fp64 column sum
        assert float((o.double() - want).abs().max()) <= 1e-5 * float(want.abs().max()) + 1e-3


def test_static_scratch_colsum_rejects_oversize_and_bad_arguments():
    x = torch.zeros(8, MAX_C + 8, dtype=torch.bfloat16, device=DEV)
    out = torch.zeros(MAX_C + 8, dtype=torch.float32, device=DEV)
    rc = _colsum_static(x, 1.0, out)
    assert rc == 1                                            # HPC_ERR_BAD_SHAPE
    msg = lib().hpc_last_error().decode()
    assert "exceeds" in msg and str(MAX_C) in msg, msg
    assert lib().hpc_colsum_static(None, 1, 8, 8, ctypes.c_float(1.0), ptr(out), stream()) == 1
    assert lib().hpc_colsum_static(ptr(x), 1, 0, 8, ctypes.c_float(1.0), ptr(out), stream()) == 1
    # the widest supported C still works after the rejections
    y = torch.ones(5, MAX_C, dtype=torch.float32, device=DEV)
    o = torch.empty(MAX_C, dtype=torch.float32, device=DEV)
    check(_colsum_static(y, 1.0, o))
    assert torch.equal(o, torch.full_like(o, 5.0))
