"""Host-side orchestration of the sm_100a kernels: layer plans, forward and hand-scheduled
backward of every block on the scale-hyperprior hot path.

Nothing here does arithmetic on tensors with torch ops: torch is used for device memory
(torch.empty / zeros), streams and autograd plumbing only; every value is produced by a
kernel of libhpcodec_sm100a.so (include/hpcodec.h).

Layout: activations are contiguous (N,H,W,C) tensors ("NHWC"); module boundaries expose
them as (N,C,H,W)-shaped views (torch channels_last), see `as_nchw` / `to_nhwc`.
"""
import ctypes
import math
import weakref
from ctypes import byref, c_double, c_float, c_int

import functools

import torch

from . import _lib
from ._lib import (BF16, DIST_GAUSS, DIST_LAPLACE, ENGINE_SIMT, ENGINE_UMMA, EPI_EXP_CLAMP, EPI_GDN, EPI_GDN_BWD,
                   EPI_IGDN, EPI_IGDN_BWD, EPI_LINEAR, EPI_RELU, EPI_RELU_MASK, EPI_SIGN_ADD, F32, check, lib, ptr,
                   stream)

# ----------------------------------------------------------------------------------------
# compute mode
# ----------------------------------------------------------------------------------------
_MODES = {
    # name: (activation dtype, igemm engine)
    "bf16": (torch.bfloat16, ENGINE_UMMA),       # product path: tcgen05/TMEM/TMA, bf16 operands, fp32 accumulate
    "fp32": (torch.float32, ENGINE_SIMT),        # exact-parity path: CUDA-core fp32 (checker for the bf16 path)
    "bf16_simt": (torch.bfloat16, ENGINE_SIMT),  # bf16 operands on CUDA cores (isolates tcgen05 bugs from rounding)
}
_mode = "bf16"


def set_compute_mode(mode):
    global _mode
    if mode not in _MODES:
        raise KeyError(f"compute mode {mode!r} not in {sorted(_MODES)}")
    _mode = mode


def get_compute_mode():
    return _mode


def lp_dtype():
    return _MODES[_mode][0]


def engine():
    return _MODES[_mode][1]


class compute_mode:
    """Context manager: `with compute_mode("fp32"): ...`"""

    def __init__(self, mode):
        self.mode = mode

    def __enter__(self):
        self.prev = get_compute_mode()
        set_compute_mode(self.mode)

    def __exit__(self, *a):
        set_compute_mode(self.prev)


# ----------------------------------------------------------------------------------------
# small utilities
# ----------------------------------------------------------------------------------------
_scratch = {}


def reduce_scratch(device):
    key = ("reduce", device)
    if key not in _scratch:
        _scratch[key] = torch.zeros(lib().hpc_reduce_scratch_bytes(), dtype=torch.uint8, device=device)
    return _scratch[key]


def sse_scratch(device, n):
    key = ("sse", device)
    need = 4096 + 4 * 64 * max(n, 1)
    if key not in _scratch or _scratch[key].numel() < need:
        _scratch[key] = torch.zeros(need, dtype=torch.uint8, device=device)
    return _scratch[key]


def to_nhwc(x, dtype=None):
    """(N,C,H,W)-shaped tensor (any strides) -> contiguous (N,H,W,C) tensor of `dtype`."""
    _lib.require_cuda(x)
    dtype = dtype or x.dtype
    v = x.permute(0, 2, 3, 1)
    if v.is_contiguous():
        if v.dtype == dtype:
            return v
        out = torch.empty(v.shape, dtype=dtype, device=x.device)
        check(lib().hpc_cast(ptr(v), _lib.dtype_code(v), ptr(out), _lib.dtype_code(out), v.numel(), stream()))
        return out
    xc = x.contiguous()
    N, C, H, W = xc.shape
    if xc.dtype != torch.float32:
        xf = torch.empty(xc.shape, dtype=torch.float32, device=x.device)
        check(lib().hpc_cast(ptr(xc), _lib.dtype_code(xc), ptr(xf), F32, xc.numel(), stream()))
        xc = xf
    out = torch.empty((N, H, W, C), dtype=dtype, device=x.device)
    check(lib().hpc_nchw_to_nhwc(ptr(xc), ptr(out), _lib.dtype_code(out), N, C, H, W, stream()))
    return out


def as_nchw(x_nhwc):
    """zero-copy (N,C,H,W) view of an NHWC tensor (== torch channels_last)."""
    return x_nhwc.permute(0, 3, 1, 2)


def cast(x, dtype):
    if x.dtype == dtype:
        return x
    out = torch.empty(x.shape, dtype=dtype, device=x.device)
    check(lib().hpc_cast(ptr(x), _lib.dtype_code(x), ptr(out), _lib.dtype_code(out), x.numel(), stream()))
    return out


def contiguous_f32(x):
    """fp32 contiguous copy/view of a tensor (module-boundary hygiene; uses the cast kernel)."""
    if x.dtype == torch.float32 and x.is_contiguous():
        return x
    if not x.is_contiguous():
        x = x.contiguous()
    return cast(x, torch.float32)


# ----------------------------------------------------------------------------------------
# tap tables
# ----------------------------------------------------------------------------------------
@functools.lru_cache(maxsize=None)
def conv_type_taps(k, s, p):
    """Conv2d(k,s,p) as one dense phase: input pixel = out*s + (kh-p)."""
    taps = [(kh - p, kw - p, kh * k + kw) for kh in range(k) for kw in range(k)]
    return taps, [(len(taps), 0, 0, 0)]


@functools.lru_cache(maxsize=None)
def deconv_type_taps(k, s, p):
    """ConvTranspose2d(k,s,p): one phase per output-pixel parity (a,b); out = i*s+a reads in[i+dh],
    dh = (a+p-kh)/s for the kh with (a+p-kh) % s == 0."""
    taps, phases = [], []
    for a in range(s):
        for b in range(s):
            begin = len(taps)
            for kh in range(k):
                if (a + p - kh) % s:
                    continue
                for kw in range(k):
                    if (b + p - kw) % s:
                        continue
                    taps.append(((a + p - kh) // s, (b + p - kw) // s, kh * k + kw))
            phases.append((len(taps) - begin, begin, a, b))
    return taps, phases


_tap_structs = {}


def _fill_taps(P, taps, phases):
    """Tap / phase tables of a parameter struct.  The ctypes arrays are built once per distinct table (a few per
    model) and block-copied afterwards: filling them element by element costs more host time than the launch."""
    key = (id(taps), id(phases), len(taps), len(phases))
    hit = _tap_structs.get(key)
    if hit is None or hit[0] is not taps or hit[1] is not phases:
        ph = (_lib.IgemmPhase * _lib.MAX_PHASES)()
        dh, dw = (_lib.c_int * _lib.MAX_TAPS)(), (_lib.c_int * _lib.MAX_TAPS)()
        for i, (nt, begin, a, b) in enumerate(phases):
            ph[i].ntaps, ph[i].tap_begin, ph[i].oh_off, ph[i].ow_off = nt, begin, a, b
        for i, (d0, d1, _) in enumerate(taps):
            dh[i], dw[i] = d0, d1
        hit = _tap_structs[key] = (taps, phases, ph, dh, dw)
    P.nphases = len(phases)
    P.phase, P.dh, P.dw = hit[2], hit[3], hit[4]


_pack_cache = {}
_capturing = False
_grad_sink = None


_grad_sink_owner = None


def set_grad_sink(sink, owner=None):
    """sink: {param.data_ptr(): zero-filled fp32 view shaped like the parameter} or None.  While armed, the
    backward functions below take the OUTPUT buffer of a parameter's gradient from the sink (each entry once), so
    the kernels write into the caller's flat gradient buffer (sharding.FlatGradients) instead of fresh tensors."""
    global _grad_sink, _grad_sink_owner
    _grad_sink = sink if sink else None
    _grad_sink_owner = owner if sink else None


def grad_sink_owner():
    return _grad_sink_owner if _grad_sink else None


_grad_ready_hook = None


def set_grad_ready_hook(fn):
    """fn(list of gradient tensors, flush) or None.  The backward schedules call it as soon as every kernel writing
    the gradients of a layer has been ENQUEUED on the current stream, so sharding.FlatGradients can start the
    all-reduce of finished slot ranges while the rest of the backward still runs."""
    global _grad_ready_hook
    _grad_ready_hook = fn


def grads_ready(grads, flush=False):
    """Called by the backward schedules (blocks_impl, autograd_ops) after every layer: the kernels writing these
    parameter gradients have all been enqueued.  flush: a parameter group ends here (start its all-reduce now)."""
    if _grad_ready_hook is not None:
        _grad_ready_hook([g for g in grads if g is not None], flush)


def grad_buffer(param, zero=False):
    """Output buffer for the gradient of `param` (fp32, param's shape)."""
    s = _grad_sink
    if s is not None and param is not None:
        v = s.pop(param.data_ptr(), None)
        if v is not None and v.shape == param.shape and v.device == param.device:
            return v                      # zero-filled by FlatGradients.zero()
    if zero:
        return torch.zeros(param.shape, dtype=torch.float32, device=param.device)
    return torch.empty(param.shape, dtype=torch.float32, device=param.device)


_captured_keys = set()


# ----------------------------------------------------------------------------------------
# side stream for work off the backward's critical path (the small-K weight gradients of h_a / h_s / the 16x16
# layers: each uses a fraction of the SMs, and so do the dgrads they would otherwise serialise with)
# ----------------------------------------------------------------------------------------
import os as _os

_side_enabled = _os.environ.get("HPC_SIDE_WGRAD", "0") != "0"      # measured r02b: 6.68 ms with, 6.60 ms without -> off
_side_state = {}


def set_side_wgrad(flag):
    global _side_enabled
    _side_enabled = bool(flag)


class side_stream:
    """`with side_stream(dev) as on: ...` runs the body on the device's side stream, ordered after everything enqueued
    on the current stream so far (fork).  `join_side_stream(dev)` makes the current stream wait for it.  Inactive (body
    runs on the current stream) when disabled or while a CUDA graph is being captured."""

    def __init__(self, device):
        self.active = _side_enabled and not _capturing
        if not self.active:
            return
        st = _side_state.get(device)
        if st is None:
            st = _side_state[device] = {"stream": torch.cuda.Stream(device=device), "fork": torch.cuda.Event(),
                                        "join": torch.cuda.Event(), "dirty": False, "held": []}
        self.st = st

    def __enter__(self):
        if self.active:
            st = self.st
            st["fork"].record()
            st["stream"].wait_event(st["fork"])
            st["dirty"] = True
            self.ctx = torch.cuda.stream(st["stream"])
            self.ctx.__enter__()
        return self.active

    def __exit__(self, *a):
        if self.active:
            self.ctx.__exit__(*a)


def hold_until_join(device, *tensors):
    """Tensors allocated on the main stream that side-stream work reads: keep them alive until the join, or the caching
    allocator may hand their storage to a later main-stream kernel while the side stream still reads it."""
    st = _side_state.get(device)
    if st is not None:
        st["held"].extend(tensors)


def join_side_stream(device):
    st = _side_state.get(device)
    if st is not None and st["dirty"]:
        st["join"].record(st["stream"])
        torch.cuda.current_stream(device).wait_event(st["join"])
        st["dirty"] = False
        st["held"].clear()


def set_capturing(flag):
    """While a CUDA graph is being captured the packed-weight cache is bypassed, so that the pack kernels are
    nodes of the graph and replays always see the current weights (operands a PackPlan re-packed INSIDE this
    capture are the exception: its launch is the graph node, and its persistent buffers are what later calls read)."""
    global _capturing
    _capturing = bool(flag)
    _captured_keys.clear()


def _plan_filled(key):
    return key in _captured_keys


def _pack_key(w, taps, rows, cols, stride_r, stride_c, dtype):
    return (id(w), rows, cols, stride_r, stride_c, dtype, tuple(t[2] for t in taps))


def _cache_get(key, w):
    """Cached operand copy of the live tensor `w` (or None).  Valid only for the very same tensor object at the same
    version and storage: ids and device addresses are both recycled after a model is freed.  The version check is
    what makes weight updates visible: torch's in-place ops bump `_version`; code that writes parameters through raw
    device pointers (solver.FusedAdamW) must bump it itself (`bump_versions`)."""
    hit = _pack_cache.get(key)
    if hit is not None and hit[0]() is w and hit[1] == w._version and hit[2] == w.data_ptr():
        return hit[3]
    return None


def _cache_put(key, w, out):
    if len(_pack_cache) > 512:
        _pack_cache.clear()
    _pack_cache[key] = (weakref.ref(w), w._version, w.data_ptr(), out)


def bump_versions(tensors):
    """Tell autograd (and every cache keyed on `_version`, the packed-operand cache above included) that `tensors`
    were modified in place by a kernel that got raw pointers."""
    torch._C._increment_version(list(tensors))


def invalidate_packs():
    """Drop every cached packed operand (e.g. after parameters were written behind torch's back)."""
    _pack_cache.clear()


class PackPlan:
    """Every packed operand copy a model's forward + backward consume -- the [tap][row][col] copies of the conv /
    transposed-conv weights for the forward and the dgrad engines, the K-/N-major copies of the 3-channel head and tail,
    the effective gamma / gamma^T / beta of the GDN layers -- rebuilt by ONE launch (hpc_pack_multi) into persistent
    buffers whenever a source parameter changed (a training step changes all of them: 33 launches otherwise).
    `ensure()` publishes the copies in the packed-operand cache, where pack_weight / pack_c3_weight / gdn_prepare find
    them.  While a CUDA graph is being captured the launch is always issued, so it becomes a node of the graph."""

    def __init__(self, dtype):
        self.dtype = dtype
        self.entries = []          # (key, sources, out object, job index)
        self.jobs = []
        self._sig = None
        self._arr = None

    def _job(self, kind, src, dst, rows, cols=0, ntaps=0, sr=0, sc=0, tap_src=(), src2=None, dst2=None, dst3=None,
             f=(0.0, 0.0, 0.0)):
        self.jobs.append(dict(kind=kind, src=src, dst=dst, rows=rows, cols=cols, ntaps=ntaps, sr=sr, sc=sc,
                              tap_src=tuple(tap_src), src2=src2, dst2=dst2, dst3=dst3, f=f))
        self._arr = None

    def add_taps(self, w, taps, rows, cols, sr, sc):
        key = _pack_key(w, taps, rows, cols, sr, sc, self.dtype)
        if any(e[0] == key for e in self.entries):
            return
        out = torch.empty((len(taps), rows, cols), dtype=self.dtype, device=w.device)
        self.entries.append((key, (w,), out))
        self._job(_lib.PACK_TAPS, w, out, rows, cols, len(taps), sr, sc, [t[2] for t in taps])

    def add_spec(self, spec, w, dgrad=True):
        self.add_taps(w, *spec.fwd_pack())
        if dgrad:
            self.add_taps(w, *spec.dgrad_pack())

    def add_c3(self, w, kind):
        Cw, Cs = w.shape[0], w.shape[1]
        out = torch.empty((Cw, 128) if kind == "k" else (80, Cw), dtype=torch.bfloat16, device=w.device)
        self.entries.append(((id(w), "c3", kind), (w,), out))
        self._job(_lib.PACK_C3_K if kind == "k" else _lib.PACK_C3_N, w, out, Cw, sr=Cs * 25, sc=25)

    def add_gdn(self, gamma_param, beta_param, beta_min=1e-6):
        C = beta_param.numel()
        ops = GdnOperands()
        ops.beta = torch.empty((C,), dtype=torch.float32, device=beta_param.device)
        both = torch.empty((2, C, C), dtype=torch.bfloat16, device=beta_param.device)
        ops.g, ops.gT = both[0], both[1]
        self.entries.append(((id(gamma_param), "gdn", float(beta_min)), (gamma_param, beta_param), ops))
        self._job(_lib.PACK_GDN, gamma_param, ops.g, C, src2=beta_param, dst2=ops.gT, dst3=ops.beta,
                  f=(GAMMA_BOUND, beta_bound(beta_min), PEDESTAL))

    def _signature(self):
        return tuple((t._version, t.data_ptr()) for e in self.entries for t in e[1])

    def ensure(self):
        if not self.jobs:
            return
        sig = self._signature()
        if sig != self._sig or _capturing:
            ptrs = tuple(s[1] for s in sig)
            if self._arr is None or self._ptrs != ptrs:
                arr = (_lib.PackJob * len(self.jobs))()
                for a, j in zip(arr, self.jobs):
                    a.kind, a.ntaps, a.rows, a.cols, a.stride_r, a.stride_c = (j["kind"], j["ntaps"], j["rows"],
                                                                              j["cols"], j["sr"], j["sc"])
                    a.src, a.src2, a.dst, a.dst2, a.dst3 = (ptr(j["src"]), ptr(j["src2"]), ptr(j["dst"]),
                                                            ptr(j["dst2"]), ptr(j["dst3"]))
                    a.f0, a.f1, a.f2 = j["f"]
                    for i, t in enumerate(j["tap_src"]):
                        a.tap_src[i] = t
                self._arr, self._ptrs = arr, ptrs
            check(lib().hpc_pack_multi(self._arr, len(self.jobs), _lib.dtype_code(self.dtype), stream()))
            self._sig = sig
            if _capturing:
                _captured_keys.update(e[0] for e in self.entries)
        elif self.entries[0][0] in _pack_cache and _pack_cache[self.entries[0][0]][3] is self.entries[0][2]:
            return                      # still published
        for key, srcs, out in self.entries:
            if len(srcs) == 1:
                _cache_put(key, srcs[0], out)
            else:        # GDN: validated on both parameters
                _pack_cache[key] = (weakref.ref(srcs[0]), tuple(t._version for t in srcs),
                                    tuple(t.data_ptr() for t in srcs), out, weakref.ref(srcs[1]))


def pack_weight(w, taps, rows, cols, stride_r, stride_c, dtype, cache=True):
    """Packed (ntaps, rows, cols) copy of a reference-layout weight in `dtype` (hpc_pack_weight).
    Cached on (tensor, version, storage) so unchanged weights are not repacked; a model's PackPlan fills the cache
    for all of its layers with one launch."""
    key = _pack_key(w, taps, rows, cols, stride_r, stride_c, dtype)
    if cache:
        hit = _cache_get(key, w)        # (during a graph capture the plan has just re-issued the pack as a graph node)
        if hit is not None and (not _capturing or _plan_filled(key)):
            return hit
        cache = not _capturing
    out = torch.empty((len(taps), rows, cols), dtype=dtype, device=w.device)
    src = (c_int * len(taps))(*[t[2] for t in taps])
    check(lib().hpc_pack_weight(ptr(w), ptr(out), _lib.dtype_code(out), len(taps), rows, cols, stride_r, stride_c,
                                src, stream()))
    if cache:
        _cache_put(key, w, out)
    return out


def unpack_wgrad(gw_packed, grad_out, taps, rows, cols, stride_r, stride_c):
    src = (c_int * len(taps))(*[t[2] for t in taps])
    check(lib().hpc_unpack_wgrad(ptr(gw_packed), ptr(grad_out), len(taps), rows, cols, stride_r, stride_c, src, 0,
                                 stream()))


# ----------------------------------------------------------------------------------------
# implicit GEMM wrappers
# ----------------------------------------------------------------------------------------
def igemm_fwd(x, w_packed, taps, phases, grid_hw, is_, os_, out_hw, Co, epi=EPI_LINEAR, bias=None, aux0=None,
              aux1=None, aux2=None, want_f32=False, want_lp=True, want_sq=False, want_abs=False, want_nrm=False):
    """x: NHWC lp tensor.  Returns dict with the requested outputs (NHWC)."""
    N, H, W, Ci = x.shape
    OH, OW = out_hw
    P = _lib.IgemmFwdParams()
    P.engine, P.dtype = engine(), _lib.dtype_code(x)
    P.x, P.w = ptr(x), ptr(w_packed)
    P.N, P.H, P.W, P.Ci = N, H, W, Ci
    P.GH, P.GW = grid_hw
    P.is_, P.os = is_, os_
    P.OH, P.OW, P.Co = OH, OW, Co
    _fill_taps(P, taps, phases)
    P.epi = epi
    P.bias, P.aux0, P.aux1, P.aux2 = ptr(bias), ptr(aux0), ptr(aux1), ptr(aux2)
    outs = {}
    shape = (N, OH, OW, Co)
    if want_f32:
        outs["f32"] = torch.empty(shape, dtype=torch.float32, device=x.device)
    if want_lp:
        outs["lp"] = torch.empty(shape, dtype=x.dtype, device=x.device)
    if want_sq:
        outs["sq"] = torch.empty(shape, dtype=x.dtype, device=x.device)
    if want_abs:
        outs["abs"] = torch.empty(shape, dtype=x.dtype, device=x.device)
    if want_nrm:
        outs["nrm"] = torch.empty(shape, dtype=x.dtype, device=x.device)
    P.out_f32, P.out_lp = ptr(outs.get("f32")), ptr(outs.get("lp"))
    P.out_sq_lp, P.out_abs_lp, P.out_nrm_lp = ptr(outs.get("sq")), ptr(outs.get("abs")), ptr(outs.get("nrm"))
    wsb = lib().hpc_igemm_fwd_workspace_bytes(byref(P))     # > 0: few-tile layer, the engine splits K over CTAs
    if wsb:
        ws = torch.empty(wsb, dtype=torch.uint8, device=x.device)
        P.workspace, P.workspace_bytes = ptr(ws), wsb
    check(lib().hpc_igemm_fwd(byref(P), stream()))
    return outs


def igemm_wgrad(lo, hi, taps, phases, hs, scale=1.0, out=None, out_strides=None):
    """Returns packed gw (ntaps, R, Cc) fp32; with `out` (a weight-shaped fp32 tensor) and
    out_strides=(stride_r, stride_c) the tcgen05 engine's reduction writes the reference layout directly."""
    N, GH, GW, R = lo.shape
    _, HH, HW, Cc = hi.shape
    P = _lib.IgemmWgradParams()
    P.engine, P.dtype = engine(), _lib.dtype_code(lo)
    P.lo, P.hi = ptr(lo), ptr(hi)
    P.N, P.GH, P.GW, P.R = N, GH, GW, R
    P.HH, P.HW, P.Cc, P.hs = HH, HW, Cc, hs
    _fill_taps(P, taps, phases)
    P.scale = scale
    if out is not None:
        assert P.engine == _lib.ENGINE_UMMA and out.is_contiguous() and out.dtype == torch.float32
        gw = out
        P.out_mode, (P.out_stride_r, P.out_stride_c) = 1, out_strides
        for i, t in enumerate(taps):
            P.out_tap_src[i] = t[2]
    else:
        gw = torch.empty((len(taps), R, Cc), dtype=torch.float32, device=lo.device)
    P.gw_f32 = ptr(gw)
    wsb = lib().hpc_igemm_wgrad_workspace_bytes(byref(P))
    ws = torch.empty(max(wsb, 1), dtype=torch.uint8, device=lo.device) if wsb else None
    P.workspace, P.workspace_bytes = ptr(ws), wsb
    check(lib().hpc_igemm_wgrad(byref(P), stream()))
    return gw


def colsum(x2d_like, scale=1.0, out=None):
    """sum over all leading dims -> (C,) fp32."""
    C = x2d_like.shape[-1]
    P = x2d_like.numel() // C
    if out is None:
        out = torch.empty((C,), dtype=torch.float32, device=x2d_like.device)
    assert out.numel() == C and out.dtype == torch.float32 and out.is_contiguous()
    nb = lib().hpc_colsum_scratch_bytes(P, C)
    sc = torch.empty(nb, dtype=torch.uint8, device=x2d_like.device)
    check(lib().hpc_colsum(ptr(x2d_like), _lib.dtype_code(x2d_like), P, C, scale, ptr(out), ptr(sc), nb, stream()))
    return out


# ----------------------------------------------------------------------------------------
# conv / deconv layers (forward, and backward = dgrad + wgrad + bias grad)
# ----------------------------------------------------------------------------------------
class ConvSpec:
    """Geometry of one nn.Conv2d / nn.ConvTranspose2d of the reference (analysis.py:55,
    synthesis.py:55-57, prior_analysis.py:54-56, prior_synthesis.py:54-56)."""

    def __init__(self, transposed, cin, cout, k, s):
        self.transposed, self.cin, self.cout, self.k, self.s = transposed, cin, cout, k, s
        self.p = k // 2
        self.K = k * k
        self.conv_taps = conv_type_taps(k, s, self.p)
        self.deconv_taps = deconv_type_taps(k, s, self.p)

    def out_hw(self, H, W):
        k, s, p = self.k, self.s, self.p
        if self.transposed:
            return (H - 1) * s - 2 * p + k + (s - 1), (W - 1) * s - 2 * p + k + (s - 1)
        return (H + 2 * p - k) // s + 1, (W + 2 * p - k) // s + 1

    # ---- packed operand geometry: (taps, rows, cols, stride_r, stride_c) of pack_weight ----------
    def fwd_pack(self):
        K, Ci, Co = self.K, self.cin, self.cout
        if self.transposed:   # weight (Ci,Co,k,k); rows=Co cols=Ci
            return self.deconv_taps[0], Co, Ci, K, Co * K
        return self.conv_taps[0], Co, Ci, Ci * K, K          # weight (Co,Ci,k,k)

    def dgrad_pack(self):
        K, Ci, Co = self.K, self.cin, self.cout
        if self.transposed:   # conv-type over the input grid; rows=Ci cols=Co from W[ci][co][k]
            return self.conv_taps[0], Ci, Co, Co * K, K
        return self.deconv_taps[0], Ci, Co, K, Ci * K        # rows=Ci cols=Co from W[co][ci][k]

    # ---- forward -------------------------------------------------------------------
    def forward(self, x, weight, bias, epi=EPI_LINEAR, **wants):
        N, H, W, Ci = x.shape
        assert Ci == self.cin, (Ci, self.cin)
        OH, OW = self.out_hw(H, W)
        wp = pack_weight(weight, *self.fwd_pack(), x.dtype)
        if self.transposed:
            taps, phases = self.deconv_taps
            return igemm_fwd(x, wp, taps, phases, (H, W), 1, self.s, (OH, OW), self.cout, epi, bias, **wants)
        taps, phases = self.conv_taps
        return igemm_fwd(x, wp, taps, phases, (OH, OW), self.s, 1, (OH, OW), self.cout, epi, bias, **wants)

    # ---- backward ------------------------------------------------------------------
    def dgrad(self, g, weight, in_hw, epi=EPI_LINEAR, **kw):
        """g: NHWC lp gradient wrt this layer's output -> gradient wrt its input."""
        N, GH_, GW_, _ = g.shape
        H, W = in_hw
        Ci = self.cin
        wp = pack_weight(weight, *self.dgrad_pack(), g.dtype)
        if self.transposed:
            taps, phases = self.conv_taps
            return igemm_fwd(g, wp, taps, phases, (H, W), self.s, 1, (H, W), Ci, epi, None, **kw)
        taps, phases = self.deconv_taps
        return igemm_fwd(g, wp, taps, phases, (GH_, GW_), 1, self.s, (H, W), Ci, epi, None, **kw)

    def wgrad(self, x, g, weight_shape_like):
        """x: layer input (NHWC lp), g: gradient wrt layer output (NHWC lp) -> grad in the
        reference weight layout (fp32)."""
        K, Ci, Co = self.K, self.cin, self.cout
        taps, phases = self.deconv_taps
        gw_ref = grad_buffer(weight_shape_like)
        lo, hi, rows, cols, sr, sc = (x, g, Ci, Co, Co * K, K) if self.transposed else (g, x, Co, Ci, Ci * K, K)
        # transposed: lo = x (rows=Ci), hi = g (cols=Co), W[ci][co][k]; else lo = g (rows=Co), hi = x, W[co][ci][k]
        if engine() == _lib.ENGINE_UMMA:
            igemm_wgrad(lo, hi, taps, phases, self.s, out=gw_ref, out_strides=(sr, sc))
        else:
            unpack_wgrad(igemm_wgrad(lo, hi, taps, phases, self.s), gw_ref, taps, rows, cols, sr, sc)
        return gw_ref


# ----------------------------------------------------------------------------------------
# GDN (modelling/layers/gdn.py:65-88)
# ----------------------------------------------------------------------------------------
_ONE_TAP = ([(0, 0, 0)], [(1, 0, 0, 0)])
PEDESTAL = float(2 ** -18) ** 2                      # gdn.py:50
GAMMA_BOUND = float((0.0 + PEDESTAL ** 2) ** 0.5)    # gdn.py:53 (pedestal squared again: SURVEY 3.5.1)


@functools.lru_cache(maxsize=None)
def beta_bound(beta_min=1e-6):
    # computed like the reference: float32 tensor arithmetic (gdn.py:52-53)
    ped = torch.tensor(PEDESTAL)
    return float((float(beta_min) + ped.clone() ** 2) ** 0.5)


def nonneg_fwd(param, bound, pedestal=PEDESTAL):
    out = torch.empty_like(param)
    check(lib().hpc_nonneg_fwd(ptr(param), ptr(out), param.numel(), bound, pedestal, stream()))
    return out


def nonneg_bwd(param, gout, bound, scale=1.0):
    g = grad_buffer(param)
    check(lib().hpc_nonneg_bwd(ptr(param), ptr(gout), ptr(g), param.numel(), bound, scale, stream()))
    return g


def square(x, relu=False):
    out = torch.empty_like(x)
    ro = torch.empty_like(x) if relu else None
    check(lib().hpc_square(ptr(x), ptr(out), _lib.dtype_code(x), x.numel(), int(relu), ptr(ro), stream()))
    return out, ro


def gdn_fused_ok(dtype, C):
    """True when the fused tcgen05 GDN kernels (gdn_umma.cu) serve this layer: they square the activation in
    shared memory, so the producing conv need not write a squared copy."""
    return engine() == ENGINE_UMMA and dtype == torch.bfloat16 and C in (64, 128, 192)


class GdnOperands:
    """Effective parameters of one GDN layer for the fused tcgen05 kernels: beta_eff (C,) fp32 and the two bf16
    operand copies of gamma_eff (gamma[i][k] for the forward, gamma^T for the backward)."""
    __slots__ = ("beta", "g", "gT")


def gdn_prepare(gamma_param, beta_param, beta_min=1e-6):
    """gdn.py:59-62 for both parameters + both operand copies in one launch (fused path only).  A model's PackPlan
    publishes them for all of its GDN layers at once; this is the stand-alone path."""
    key = (id(gamma_param), "gdn", float(beta_min))
    hit = _pack_cache.get(key)
    if (hit is not None and len(hit) == 5 and hit[0]() is gamma_param and hit[4]() is beta_param
            and hit[1] == (gamma_param._version, beta_param._version)
            and hit[2] == (gamma_param.data_ptr(), beta_param.data_ptr()) and (not _capturing or _plan_filled(key))):
        return hit[3]
    C = beta_param.numel()
    ops = GdnOperands()
    ops.beta = torch.empty((C,), dtype=torch.float32, device=beta_param.device)
    ops.g = torch.empty((2, C, C), dtype=torch.bfloat16, device=beta_param.device)
    ops.gT = ops.g[1]
    check(lib().hpc_gdn_prepare(ptr(gamma_param), ptr(beta_param), C, GAMMA_BOUND, beta_bound(beta_min), PEDESTAL,
                                ptr(ops.beta), ptr(ops.g[0]), ptr(ops.gT), stream()))
    ops.g = ops.g[0]
    return ops


def gdn_forward(x, xsq, gamma_eff, beta_eff, inverse=False, want_nrm=True):
    """x (, xsq): NHWC lp; gamma_eff (C,C,1,1) fp32, beta_eff (C,) fp32 -> (y, nrm) NHWC lp.
    xsq may be None (computed here, or never materialised on the fused path).  gamma_eff may be a GdnOperands."""
    N, H, W, C = x.shape
    if gdn_fused_ok(x.dtype, C):
        if isinstance(gamma_eff, GdnOperands):
            wp, beta_eff = gamma_eff.g, gamma_eff.beta
        else:
            wp = pack_weight(gamma_eff, _ONE_TAP[0], C, C, C, 1, x.dtype, cache=False)
        y, nrm = torch.empty_like(x), (torch.empty_like(x) if want_nrm else None)   # no nrm: no backward will follow
        check(lib().hpc_gdn_fwd_umma(ptr(x), ptr(wp), ptr(beta_eff), N * H * W, C, int(inverse), ptr(y),
                                     ptr(nrm) if want_nrm else None, stream()))
        return y, nrm
    if xsq is None:
        xsq, _ = square(x)
    wp = pack_weight(gamma_eff, _ONE_TAP[0], C, C, C, 1, x.dtype, cache=False)
    o = igemm_fwd(xsq, wp, _ONE_TAP[0], _ONE_TAP[1], (H, W), 1, 1, (H, W), C, EPI_IGDN if inverse else EPI_GDN,
                  beta_eff, aux0=x, want_nrm=True)
    return o["lp"], o["nrm"]


# Measured (r02i, one box visit, 64 x 256^2 training step): saved r 6.70 ms / gdn class 1.88 ms; recompute 6.73 ms / 1.98 ms.
# The backward kernel is bound by instruction issue and latency (ncu: 210 M warp instructions in 420 us at the 128^2 layer,
# issue slots 51 % busy, 19 warps resident), not by HBM, so two HBM passes fewer do not pay for ~25 % more instructions.
# Off by default; on, it saves the r tensors (~1 GB of saved activations at batch 64).
_gdn_recompute = _os.environ.get("HPC_GDN_RECOMPUTE", "0") != "0"


def set_gdn_recompute(flag):
    global _gdn_recompute
    _gdn_recompute = bool(flag)


def gdn_bwd_recomputes(dtype, C, gamma_eff):
    """True when the backward of this GDN layer rebuilds r = n^-1/2 itself (hpc_gdn_bwd_fused_rc_umma): the forward
    then need not write r and the backward need not read it (5 HBM passes per layer and step instead of 7).
    Opt-in with HPC_GDN_RECOMPUTE=1 or hotpath.set_gdn_recompute(True)."""
    return (_gdn_recompute and gdn_fused_ok(dtype, C) and C in (128, 192) and isinstance(gamma_eff, GdnOperands)
            and not _capturing)


def gdn_bwd_has_dxsum(x):
    """True when gdn_backward(x, ..., want_dxsum=True) will return the column sums of dx."""
    return gdn_fused_ok(x.dtype, x.shape[-1]) and x.shape[-1] in (128, 192)


def gdn_backward(g, x, xsq, nrm, gamma_eff, inverse=False, need_dx=True, want_dxsum=False, dxsum_out=None,
                 reparam=None):
    """SURVEY A.4.  Returns (dx lp, dgamma_eff (C,C,1,1) fp32, dbeta_eff (C,) fp32) and, with want_dxsum, a
    fourth value: the column sums of dx (the bias gradient of the layer that feeds this GDN) when the fused
    kernel produced them, else None.  reparam = (gamma_param, beta_param, beta_min): return the gradients of the
    RAW parameters instead (gdn.py:42-62 backward), fused into the finishing kernel where that exists."""
    out = _gdn_backward(g, x, xsq, nrm, gamma_eff, inverse, need_dx, dxsum_out if want_dxsum else None, reparam)
    return out if want_dxsum else out[:3]


def _gdn_backward(g, x, xsq, nrm, gamma_eff, inverse, need_dx, dxsum_out=None, reparam=None):
    dx, dgamma, dbeta, dxsum, done = _gdn_backward_eff(g, x, xsq, nrm, gamma_eff, inverse, need_dx, dxsum_out, reparam)
    if reparam is not None and not done:
        gp, bp, beta_min = reparam
        dgamma = nonneg_bwd(gp, dgamma, GAMMA_BOUND)
        dbeta = nonneg_bwd(bp, dbeta, beta_bound(beta_min))
    return dx, dgamma, dbeta, dxsum


def _gdn_backward_eff(g, x, xsq, nrm, gamma_eff, inverse, need_dx, dxsum_out, reparam):
    N, H, W, C = x.shape
    sgn = 0.5 if inverse else -0.5
    if gdn_fused_ok(x.dtype, C) and need_dx:
        P = N * H * W
        if isinstance(gamma_eff, GdnOperands):
            wp = gamma_eff.gT
        else:
            wp = pack_weight(gamma_eff, _ONE_TAP[0], C, C, 1, C, x.dtype, cache=False)   # rows=k, cols=i: gamma[i][k]
        if C in (128, 192):      # d(gamma), d(beta), the dx column sums and the reparametrisation in two launches
            dx = torch.empty_like(x)
            dev = x.device
            if reparam is not None:
                gp, bp, beta_min = reparam
                dgamma, dbeta = grad_buffer(gp), grad_buffer(bp)
                gpp, gb, bpp, bb = ptr(gp), GAMMA_BOUND, ptr(bp), beta_bound(beta_min)
            else:
                dgamma = torch.empty((C, C, 1, 1), dtype=torch.float32, device=dev)
                dbeta = torch.empty((C,), dtype=torch.float32, device=dev)
                gpp, gb, bpp, bb = None, 0.0, None, 0.0
            dxsum = dxsum_out if dxsum_out is not None else torch.empty((C,), dtype=torch.float32, device=dev)
            wsb = lib().hpc_gdn_bwd_fused_umma_workspace_bytes(P, C)
            ws = torch.empty(wsb, dtype=torch.uint8, device=dev)
            if nrm is None:       # the forward kept no r: the kernel rebuilds it from x, gamma and beta
                if not isinstance(gamma_eff, GdnOperands):
                    raise RuntimeError("gdn_backward: r was not saved and the effective beta is not at hand")
                check(lib().hpc_gdn_bwd_fused_rc_umma(ptr(g), ptr(x), ptr(gamma_eff.beta), ptr(wp), P, C, int(inverse),
                                                      ptr(dx), gpp, gb, bpp, bb, ptr(dgamma), ptr(dbeta), ptr(dxsum),
                                                      ptr(ws), wsb, stream()))
            else:
                check(lib().hpc_gdn_bwd_fused_umma(ptr(g), ptr(x), ptr(nrm), ptr(wp), P, C, int(inverse), ptr(dx), gpp, gb,
                                                   bpp, bb, ptr(dgamma), ptr(dbeta), ptr(dxsum), ptr(ws), wsb, stream()))
            return dx, dgamma, dbeta, dxsum, True
        dx, t, sq = torch.empty_like(x), torch.empty_like(x), torch.empty_like(x)
        parts = torch.empty((lib().hpc_gdn_bwd_umma_parts(P), C), dtype=torch.float32, device=x.device)
        check(lib().hpc_gdn_bwd_umma(ptr(g), ptr(x), ptr(nrm), ptr(wp), P, C, int(inverse), ptr(dx), ptr(t), ptr(sq),
                                     ptr(parts), stream()))
        dgamma = igemm_wgrad(t, sq, _ONE_TAP[0], _ONE_TAP[1], 1, sgn).view(C, C, 1, 1)
        return dx, dgamma, colsum(parts, sgn), None, False
    if xsq is None:
        xsq, _ = square(x)
    t = torch.empty_like(x)
    check(lib().hpc_gdn_bwd_t(ptr(g), ptr(x), ptr(nrm), ptr(t), _lib.dtype_code(x), x.numel(), int(inverse),
                              stream()))
    dx = None
    if need_dx:
        wp = pack_weight(gamma_eff, _ONE_TAP[0], C, C, 1, C, x.dtype, cache=False)  # rows=k, cols=i: gamma[i][k]
        dx = igemm_fwd(t, wp, _ONE_TAP[0], _ONE_TAP[1], (H, W), 1, 1, (H, W), C,
                       EPI_IGDN_BWD if inverse else EPI_GDN_BWD, None, aux0=g, aux1=x, aux2=nrm)["lp"]
    dgamma = igemm_wgrad(t, xsq, _ONE_TAP[0], _ONE_TAP[1], 1, sgn).view(C, C, 1, 1)
    dbeta = colsum(t, sgn)
    return dx, dgamma, dbeta, None, False


# ----------------------------------------------------------------------------------------
# thin-channel layers
# ----------------------------------------------------------------------------------------
def _thin_umma_ok(Cw, Cs, dt):
    return engine() == ENGINE_UMMA and dt == torch.bfloat16 and Cs == 3 and Cw in (64, 128, 192, 256)


def pack_c3_weight(weight, kind):
    """bf16 operand of the tcgen05 thin kernels from a (Cw,3,5,5)-laid-out weight (Conv2d(3->Cw) weight or
    ConvTranspose2d(Cw->3) weight): kind "k" -> [Cw][128] (k contiguous), "n" -> [80][Cw].  Cached like pack_weight."""
    key = (id(weight), "c3", kind)
    hit = _cache_get(key, weight)
    if hit is not None and (not _capturing or _plan_filled(key)):
        return hit
    Cw, Cs = weight.shape[0], weight.shape[1]
    out = torch.empty((Cw, 128) if kind == "k" else (80, Cw), dtype=torch.bfloat16, device=weight.device)
    check(lib().hpc_pack_c3_weight(ptr(weight), Cs * 25, 25, Cw, ptr(out) if kind == "k" else None,
                                   ptr(out) if kind == "n" else None, stream()))
    if not _capturing:
        _cache_put(key, weight, out)
    return out


def conv_c3_forward(x_nchw, weight, bias, want_sq=True, want_f32=False):
    """Conv2d(3->C,5,2,2) on an NCHW fp32 image -> NHWC lp (+ squares)."""
    N, Cs, H, W = x_nchw.shape
    Cw = weight.shape[0]
    dt = lp_dtype()
    shape = (N, H // 2, W // 2, Cw)
    out = torch.empty(shape, dtype=dt, device=x_nchw.device)
    sq = torch.empty(shape, dtype=dt, device=x_nchw.device) if want_sq else None
    # (the tcgen05 kernel fetches its input patches by TMA: rows of whole 16-byte units, 16-byte aligned base)
    if _thin_umma_ok(Cw, Cs, dt) and Cw <= 192 and not want_f32 and W % 4 == 0 and x_nchw.data_ptr() % 16 == 0:
        check(lib().hpc_conv_c3_fwd_umma(ptr(x_nchw), ptr(pack_c3_weight(weight, "k")), ptr(bias), N, H, W, Cw, ptr(out),
                                         ptr(sq), stream()))
        return out, sq, None
    of = torch.empty(shape, dtype=torch.float32, device=x_nchw.device) if want_f32 else None
    check(lib().hpc_conv_c3_fwd(ptr(x_nchw), ptr(weight), Cs * 25, 25, ptr(bias), N, Cs, H, W, Cw, ptr(of), ptr(out),
                                ptr(sq), _lib.dtype_code(dt), stream()))
    return out, sq, of


def wgrad_c3(wide_nhwc, thin_nchw, Cw, Cs, weight=None):
    """gw (Cw,Cs,5,5) = sum wide[n,i,j,cw]*thin[n,cs,2i-2+kh,2j-2+kw].  `weight`: the parameter this is the
    gradient of (selects the output buffer, see grad_buffer)."""
    N, H, W, _ = wide_nhwc.shape
    if weight is not None and tuple(weight.shape) == (Cw, Cs, 5, 5):
        gw = grad_buffer(weight)
    else:
        gw = torch.empty((Cw, Cs, 5, 5), dtype=torch.float32, device=wide_nhwc.device)
    # (the tcgen05 kernel fetches the image patches by TMA: rows of 2W floats in whole 16-byte units, aligned base)
    if _thin_umma_ok(Cw, Cs, wide_nhwc.dtype) and W % 2 == 0 and thin_nchw.data_ptr() % 16 == 0:
        nb = lib().hpc_wgrad_c3_umma_workspace_bytes(N, Cw, H, W)
        sc = torch.empty(nb, dtype=torch.uint8, device=wide_nhwc.device)
        check(lib().hpc_wgrad_c3_umma(ptr(wide_nhwc), ptr(thin_nchw), N, Cw, H, W, ptr(gw), Cs * 25, 25, ptr(sc), nb,
                                      stream()))
        return gw
    nb = lib().hpc_wgrad_c3_scratch_bytes(N, Cw, Cs, H, W)
    sc = torch.empty(nb, dtype=torch.uint8, device=wide_nhwc.device)
    check(lib().hpc_wgrad_c3(ptr(wide_nhwc), _lib.dtype_code(wide_nhwc), ptr(thin_nchw), N, Cw, Cs, H, W, ptr(gw),
                             Cs * 25, 25, ptr(sc), nb, stream()))
    return gw


def deconv_c3_forward(x_nhwc, weight, bias):
    """ConvTranspose2d(C->3,5,2,2,1): NHWC lp -> NCHW fp32 image."""
    N, H, W, Cw = x_nhwc.shape
    Cs = weight.shape[1]
    out = torch.empty((N, Cs, 2 * H, 2 * W), dtype=torch.float32, device=x_nhwc.device)
    if _thin_umma_ok(Cw, Cs, x_nhwc.dtype) and Cw <= 192:
        check(lib().hpc_deconv_c3_fwd_umma(ptr(x_nhwc), ptr(pack_c3_weight(weight, "n")), ptr(bias), N, Cw, H, W,
                                           ptr(out), stream()))
        return out
    check(lib().hpc_deconv_c3_fwd(ptr(x_nhwc), _lib.dtype_code(x_nhwc), ptr(weight), Cs * 25, 25, ptr(bias), N, Cw, H,
                                  W, Cs, ptr(out), stream()))
    return out


def deconv_c3_backward(g_nchw, x_nhwc, weight, need_dx=True, bias=None):
    """g: (N,3,2H,2W) fp32.  dgrad = Conv2d(3->C) with the same weight layout; wgrad; bias grad."""
    N, Cs, H2, W2 = g_nchw.shape
    Cw = weight.shape[0]
    dt = x_nhwc.dtype
    dx = None
    if need_dx:
        dx = torch.empty((N, H2 // 2, W2 // 2, Cw), dtype=dt, device=g_nchw.device)
        if _thin_umma_ok(Cw, Cs, dt) and Cw <= 192 and W2 % 4 == 0 and g_nchw.data_ptr() % 16 == 0:
            check(lib().hpc_conv_c3_fwd_umma(ptr(g_nchw), ptr(pack_c3_weight(weight, "k")), None, N, H2, W2, Cw, ptr(dx),
                                             None, stream()))
        else:
            check(lib().hpc_conv_c3_fwd(ptr(g_nchw), ptr(weight), Cs * 25, 25, None, N, Cs, H2, W2, Cw, None, ptr(dx),
                                        None, _lib.dtype_code(dt), stream()))
    gw = wgrad_c3(x_nhwc, g_nchw, Cw, Cs, weight)
    gb = grad_buffer(bias) if bias is not None else torch.empty((Cs,), dtype=torch.float32, device=g_nchw.device)
    sc = torch.empty((N * Cs,), dtype=torch.float32, device=g_nchw.device)
    check(lib().hpc_nchw_channel_sum(ptr(g_nchw), N, Cs, H2 * W2, ptr(gb), ptr(sc), stream()))
    return dx, gw, gb


# ----------------------------------------------------------------------------------------
# entropy models
# ----------------------------------------------------------------------------------------
def _cdf_struct(tensors):
    """tensors: dict with w0..w3, b0..b3, f0..f2 (reference layouts, fp32 contiguous)."""
    s = _lib.CdfParams()
    for i in range(4):
        s.w[i] = ptr(tensors[f"w{i}"])
        s.b[i] = ptr(tensors[f"b{i}"])
    for i in range(3):
        s.f[i] = ptr(tensors[f"f{i}"])
    return s


def _check_cdf_dims(t):
    ok = (t["w0"].shape[2:] == (3, 1) and t["w1"].shape[2:] == (3, 3) and t["w2"].shape[2:] == (3, 3)
          and t["w3"].shape[2:] == (1, 3))
    if not ok:
        raise NotImplementedError("the factorised-prior kernel is specialised for ENTROPY_MODEL.DIMS == [3,3,3] "
                                  "(utils/config.py:66); got " + str([tuple(t[f'w{i}'].shape) for i in range(4)]))


def _rank_salt():
    """Ranks of a data-parallel job are usually seeded identically (the reference seeds every process with cfg.SEED,
    utils/deterministic.py:21-29): mix the rank into the Philox key so replicas draw different quantisation noise."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized():
        return (dist.get_rank() * 0x9E3779B97F4A7C15) & (2 ** 62 - 1)
    return 0


def next_seed():
    """(seed, offset) for the in-kernel Philox stream, drawn from torch's CPU generator so that
    torch.manual_seed() makes runs reproducible (no device sync); the rank is mixed into the seed."""
    v = torch.randint(0, 2 ** 62, (2,), dtype=torch.int64)
    return int(v[0]) ^ _rank_salt(), int(v[1])


def factorized_forward(z, cdf, training, noise=None, bin_=1.0, want_probs=True, want_lp=False):
    """z: NHWC fp32 (N,h,w,C).  Returns z_tilde f32, z_tilde lp|None, probs|None, bits (1,)."""
    _check_cdf_dims(cdf)
    C = z.shape[-1]
    P = z.numel() // C
    zt = torch.empty_like(z)
    ztl = torch.empty(z.shape, dtype=lp_dtype(), device=z.device) if want_lp else None
    probs = torch.empty_like(z) if want_probs else None
    bits = torch.empty((1,), dtype=torch.float32, device=z.device)
    seed, off = (0, 0) if (noise is not None or not training) else next_seed()
    s = _cdf_struct(cdf)
    check(lib().hpc_factorized_fwd(ptr(z), ptr(noise), byref(s), P, C, int(training), bin_, seed, off, ptr(zt),
                                   ptr(ztl), _lib.dtype_code(lp_dtype()), ptr(probs), ptr(bits),
                                   ptr(reduce_scratch(z.device)), stream()))
    return zt, ztl, probs, bits


def factorized_backward(zt, cdf, training, g_bits, g_scale, g_probs=None, g_zt=None, bin_=1.0):
    """Returns (gz fp32, dict of parameter grads in the reference layouts)."""
    C = zt.shape[-1]
    P = zt.numel() // C
    gz = torch.empty_like(zt)
    grads = {k: grad_buffer(v, zero=True) for k, v in cdf.items()}
    s, gs = _cdf_struct(cdf), _cdf_struct(grads)
    nb = lib().hpc_factorized_bwd_scratch_bytes(P, C)      # per-row-block partials: fixed-order (reproducible) sums
    part = torch.empty(nb // 4, dtype=torch.float32, device=zt.device)
    check(lib().hpc_factorized_bwd(ptr(zt), byref(s), P, C, int(training), bin_, ptr(g_bits), g_scale, ptr(g_probs),
                                   ptr(g_zt), ptr(gz), byref(gs), ptr(part), nb, stream()))
    return gz, grads


def conditional_forward(y, scale, dist, training, noise=None, bin_=1.0, mean=0.0, want_probs=True, want_lp=False):
    n = y.numel()
    yt = torch.empty_like(y)
    ytl = torch.empty(y.shape, dtype=lp_dtype(), device=y.device) if want_lp else None
    probs = torch.empty_like(y) if want_probs else None
    bits = torch.empty((1,), dtype=torch.float32, device=y.device)
    seed, off = (0, 0) if (noise is not None or not training) else next_seed()
    check(lib().hpc_conditional_fwd(ptr(y), ptr(scale), ptr(noise), n, dist, int(training), bin_, mean, seed, off,
                                    ptr(yt), ptr(ytl), _lib.dtype_code(lp_dtype()), ptr(probs), ptr(bits),
                                    ptr(reduce_scratch(y.device)), stream()))
    return yt, ytl, probs, bits


def conditional_backward(yt, scale, dist, training, g_bits, g_scale, g_probs=None, g_yt=None, fuse_exp_clamp=False,
                         want_gs_f32=True, want_gs_lp=False, bin_=1.0, mean=0.0):
    n = yt.numel()
    gy = torch.empty_like(yt)
    gs = torch.empty_like(scale) if want_gs_f32 else None
    gsl = torch.empty(scale.shape, dtype=lp_dtype(), device=yt.device) if want_gs_lp else None
    check(lib().hpc_conditional_bwd(ptr(yt), ptr(scale), n, dist, int(training), bin_, mean, ptr(g_bits), g_scale,
                                    ptr(g_probs), ptr(g_yt), _lib.dtype_code(g_yt) if g_yt is not None else F32,
                                    int(fuse_exp_clamp), ptr(gy), ptr(gs), ptr(gsl), _lib.dtype_code(lp_dtype()),
                                    stream()))
    return gy, gs, gsl


# ----------------------------------------------------------------------------------------
# distortion
# ----------------------------------------------------------------------------------------
def clamp01_sse(x, raw, scale=1.0, do_clamp=True, want_xt=True):
    """x, raw: (N,...) fp32 contiguous.  Returns (x_tilde|None, sse_per_image (N,))."""
    N = x.shape[0]
    per = x.numel() // N
    xt = torch.empty_like(raw) if want_xt else None
    sse = torch.empty((N,), dtype=torch.float32, device=x.device)
    check(lib().hpc_clamp01_sse_fwd(ptr(x), ptr(raw), ptr(xt), N, per, scale, int(do_clamp), ptr(sse),
                                    ptr(sse_scratch(x.device, N)), stream()))
    return xt, sse


def clamp01_backward(raw, x, g_xt, mse_coef, mse_coef_dev=None, do_clamp=True):
    g = torch.empty_like(raw)
    check(lib().hpc_clamp01_bwd(ptr(raw), ptr(x), ptr(g_xt), raw.numel(), mse_coef, ptr(mse_coef_dev), int(do_clamp),
                                ptr(g), stream()))
    return g


def relu_mask(g, x):
    out = torch.empty_like(g)
    check(lib().hpc_relu_mask(ptr(g), ptr(x), ptr(out), _lib.dtype_code(g), g.numel(), stream()))
    return out


_filters = {}


def gauss_filter_1d(size, sigma, device):
    """1-D factor g of the reference window softmax2D(-(x^2+y^2)/2s^2) = g (x) g  (loss.py:113-121,
    SURVEY A.5).  Tiny host computation, cached per device."""
    key = (size, sigma, device)
    if key not in _filters:
        r = [i + 0.5 - size / 2 for i in range(size)]
        e = [math.exp(-(v * v) / (2.0 * sigma * sigma)) for v in r]
        s = sum(e)
        _filters[key] = torch.tensor([v / s for v in e], dtype=torch.float32).to(device)
    return _filters[key]


class SsimPyramid:
    """Forward state of an L-level (MS-)SSIM evaluation (K7)."""

    def __init__(self, img1, img2, levels, fs, sigma, c1, c2, in_scale):
        N, C, H, W = img1.shape
        self.N, self.C, self.levels, self.fs, self.c1, self.c2, self.in_scale = N, C, levels, fs, c1, c2, in_scale
        self.filt = gauss_filter_1d(fs, sigma, img1.device)
        self.imgs = [(img1, img2, H, W, in_scale)]
        self.sums = torch.empty((levels, 2, N), dtype=torch.float32, device=img1.device)
        self.counts = []
        fx = torch.empty((levels, 2 * N), dtype=torch.int64, device=img1.device)   # fixed-point accumulators
        a, b, h, w, sc = img1, img2, H, W, in_scale
        for l in range(levels):
            if h < fs or w < fs:
                raise ValueError(f"(MS-)SSIM level {l}: image {h}x{w} smaller than the {fs}-tap window "
                                 "(the reference's F.conv2d raises here too)")
            check(lib().hpc_ssim_level_fwd(ptr(a), ptr(b), N, C, h, w, ptr(self.filt), fs, sc, c1, c2,
                                           ptr(self.sums[l, 0]), ptr(self.sums[l, 1]), ptr(fx[l]), stream()))
            self.counts.append(float(C * (h - fs + 1) * (w - fs + 1)))
            if l < levels - 1:
                oh, ow = (h + 1) // 2, (w + 1) // 2
                a2 = torch.empty((N, C, oh, ow), dtype=torch.float32, device=img1.device)
                b2 = torch.empty((N, C, oh, ow), dtype=torch.float32, device=img1.device)
                # level 0 is stored unscaled; the pooled level carries the x255 scale from here on
                check(lib().hpc_avgpool2_fwd(ptr(a), ptr(a2), N * C, h, w, sc, stream()))
                check(lib().hpc_avgpool2_fwd(ptr(b), ptr(b2), N * C, h, w, sc, stream()))
                a, b, h, w, sc = a2, b2, oh, ow, 1.0
                self.imgs.append((a, b, h, w, sc))

    def combine(self, kind, flags, eps, weights):
        L, N = self.levels, self.N
        n_out = N if kind in (_lib.MSSSIM_METRIC, _lib.SSIM_METRIC) or (kind == _lib.SSIM_LOSS and flags & 1) else 1
        out = torch.empty((n_out,), dtype=torch.float32, device=self.sums.device)
        cnt = (c_double * L)(*self.counts)
        wts = (c_float * L)(*[float(w) for w in weights])
        check(lib().hpc_msssim_combine_fwd(ptr(self.sums), cnt, L, N, wts, kind, flags, eps, ptr(out), stream()))
        return out

    def backward(self, kind, flags, eps, weights, g_out, g_scale=1.0):
        """gradient wrt img2 (level-0, unscaled units); upstream = g_out (device) * g_scale."""
        L, N, C = self.levels, self.N, self.C
        g_sums = torch.empty_like(self.sums)
        cnt = (c_double * L)(*self.counts)
        wts = (c_float * L)(*[float(w) for w in weights])
        check(lib().hpc_msssim_combine_bwd(ptr(self.sums), cnt, L, N, wts, kind, flags, eps, ptr(g_out), g_scale, ptr(g_sums),
                                           stream()))
        g_next = None  # gradient wrt the images of level l+1
        for l in range(L - 1, -1, -1):
            a, b, h, w, sc = self.imgs[l]
            oh, ow = h - self.fs + 1, w - self.fs + 1
            g = torch.empty((N, C, h, w), dtype=torch.float32, device=a.device)
            acc = 0
            if g_next is not None:
                # pooled level = sc_l * avgpool(level l)  -> scatter back
                check(lib().hpc_avgpool2_bwd(ptr(g_next), ptr(g), N * C, h, w, sc, 0, stream()))
                acc = 1
            maps = torch.empty((3, N * C, oh, ow), dtype=torch.float32, device=a.device)
            check(lib().hpc_ssim_level_bwd(ptr(a), ptr(b), N, C, h, w, ptr(self.filt), self.fs, sc, self.c1, self.c2,
                                           ptr(g_sums[l, 0]), ptr(g_sums[l, 1]), ptr(maps), ptr(g), acc, stream()))
            g_next = g
        return g_next


def psnr_from_sse(sse, per_image, max_val):
    out = torch.empty_like(sse)
    check(lib().hpc_psnr_from_sse(ptr(sse), sse.numel(), per_image, max_val, ptr(out), stream()))
    return out


def rd_combine(z_bits, y_bits, terms, lmbda, inv_num_pixels):
    """terms: list of (tensor, scale).  Returns (8,) fp32: z_entropy, y_entropy, bpp, total, d0..d3."""
    out = torch.empty((8,), dtype=torch.float32, device=z_bits.device)
    T = _lib.DistTerms()
    T.n_terms = len(terms)
    for i, (t, sc) in enumerate(terms):
        T.dist[i], T.n[i], T.scale[i] = ptr(t), t.numel(), sc
    check(lib().hpc_rd_combine(ptr(z_bits), ptr(y_bits), byref(T), lmbda, inv_num_pixels, ptr(out), stream()))
    return out


def exp_clamp_backward(sigma, g_sigma):
    out = torch.empty(sigma.shape, dtype=lp_dtype(), device=sigma.device)
    check(lib().hpc_exp_clamp_bwd(ptr(sigma), ptr(g_sigma), ptr(out), _lib.dtype_code(out), sigma.numel(), stream()))
    return out
