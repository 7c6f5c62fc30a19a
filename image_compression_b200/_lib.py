"""ctypes binding of libhpcodec_sm100a.so (include/hpcodec.h) and its in-tree build.

There is no CPU fallback: every op in this package goes through the C ABI; if the shared
library is missing the import of any op raises with the build command to run.
"""
import ctypes
import os
import subprocess
import sys
from ctypes import (POINTER, Structure, byref, c_char_p, c_double, c_float, c_int, c_int64, c_size_t,
                    c_uint64, c_void_p)

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(_HERE, "csrc")
LIB_PATH = os.path.join(_HERE, "libhpcodec_sm100a.so")
SOURCES = ["hpc_api.cu", "rans_kernels.cu", "entropy_kernels.cu", "loss_kernels.cu", "igemm_simt.cu", "thin_kernels.cu",
           "igemm_umma.cu", "thin_umma.cu", "gdn_umma.cu", "misc_kernels.cu", "optim_kernels.cu"]
# no --use_fast_math: IEEE division / sqrt / expf are part of the parity contract
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC"]

F32, BF16 = 0, 1
ENGINE_SIMT, ENGINE_UMMA = 0, 1
MAX_TAPS, MAX_PHASES = 25, 4

(EPI_LINEAR, EPI_RELU, EPI_EXP_CLAMP, EPI_GDN, EPI_IGDN, EPI_GDN_BWD, EPI_IGDN_BWD, EPI_RELU_MASK,
 EPI_SIGN_ADD) = range(9)
DIST_LAPLACE, DIST_GAUSS = 0, 1
(PROF_IGEMM_FWD, PROF_IGEMM_WGRAD, PROF_CONDITIONAL, PROF_FACTORIZED, PROF_THIN, PROF_DISTORTION, PROF_GDN) = range(7)
PROF_NAMES = ("igemm_fwd_umma", "igemm_wgrad_umma", "conditional", "factorized", "thin_c3", "distortion", "gdn_fused",
              "optim")
MSSSIM_LOSS_LOG, MSSSIM_LOSS_PLAIN, MSSSIM_METRIC, SSIM_LOSS, SSIM_METRIC = range(5)


def _needs_build():
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + [os.path.join(_HERE, "..", "include", "hpcodec.h")]
    return any(os.path.getmtime(d) > t for d in deps)


def build_library(force=False, verbose=False):
    """nvcc-compile csrc/*.cu for sm_100a into libhpcodec_sm100a.so next to this file."""
    if not force and not _needs_build():
        return LIB_PATH
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    objdir = os.path.join(_HERE, "build")
    os.makedirs(objdir, exist_ok=True)
    procs = []
    for src in SOURCES:
        obj = os.path.join(objdir, src.replace(".cu", ".o"))
        srcp = os.path.join(CSRC, src)
        if not force and os.path.exists(obj) and os.path.getmtime(obj) > max(
                os.path.getmtime(os.path.join(CSRC, f)) for f in os.listdir(CSRC)
                if f == src or f.endswith(".cuh")) and os.path.getmtime(obj) > os.path.getmtime(
                    os.path.join(_HERE, "..", "include", "hpcodec.h")):
            continue
        cmd = [nvcc] + NVCC_FLAGS + ["-c", srcp, "-o", obj]
        if verbose:
            print(" ".join(cmd), file=sys.stderr)
        procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)))
    for src, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0:
            raise RuntimeError(f"nvcc failed for {src}:\n{out.decode()}")
        if verbose and out:
            print(out.decode(), file=sys.stderr)
    objs = [os.path.join(objdir, s.replace(".cu", ".o")) for s in SOURCES]
    cmd = [nvcc, "-shared", "-o", LIB_PATH] + objs + ["-gencode", "arch=compute_100a,code=sm_100a"]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout.decode()}")
    return LIB_PATH


# ----------------------------------------------------------------------------------------
# struct mirrors of include/hpcodec.h
# ----------------------------------------------------------------------------------------
class CdfParams(Structure):
    _fields_ = [("w", c_void_p * 4), ("b", c_void_p * 4), ("f", c_void_p * 3)]


class IgemmPhase(Structure):
    _fields_ = [("ntaps", c_int), ("tap_begin", c_int), ("oh_off", c_int), ("ow_off", c_int)]


class IgemmFwdParams(Structure):
    _fields_ = [
        ("engine", c_int), ("dtype", c_int), ("x", c_void_p), ("w", c_void_p),
        ("N", c_int), ("H", c_int), ("W", c_int), ("Ci", c_int),
        ("GH", c_int), ("GW", c_int), ("is_", c_int), ("os", c_int),
        ("OH", c_int), ("OW", c_int), ("Co", c_int), ("nphases", c_int),
        ("phase", IgemmPhase * MAX_PHASES), ("dh", c_int * MAX_TAPS), ("dw", c_int * MAX_TAPS),
        ("epi", c_int), ("bias", c_void_p), ("aux0", c_void_p), ("aux1", c_void_p), ("aux2", c_void_p),
        ("out_f32", c_void_p), ("out_lp", c_void_p), ("out_sq_lp", c_void_p), ("out_abs_lp", c_void_p),
        ("out_nrm_lp", c_void_p), ("workspace", c_void_p), ("workspace_bytes", c_size_t),
    ]


class IgemmWgradParams(Structure):
    _fields_ = [
        ("engine", c_int), ("dtype", c_int), ("lo", c_void_p), ("hi", c_void_p),
        ("N", c_int), ("GH", c_int), ("GW", c_int), ("R", c_int),
        ("HH", c_int), ("HW", c_int), ("Cc", c_int), ("hs", c_int), ("nphases", c_int),
        ("phase", IgemmPhase * MAX_PHASES), ("dh", c_int * MAX_TAPS), ("dw", c_int * MAX_TAPS),
        ("scale", c_float), ("gw_f32", c_void_p), ("workspace", c_void_p), ("workspace_bytes", c_size_t),
        ("out_mode", c_int), ("out_accumulate", c_int), ("out_stride_r", c_int64), ("out_stride_c", c_int64),
        ("out_tap_src", c_int * MAX_TAPS),
    ]


class AdamwSeg(Structure):
    _fields_ = [("p", c_void_p), ("g", c_void_p), ("m", c_void_p), ("v", c_void_p), ("n", c_int64),
                ("decay", c_float), ("neg_step_size", c_float), ("bc2_sqrt", c_float), ("reserved", c_float)]


class PackJob(Structure):
    _fields_ = [("kind", c_int), ("ntaps", c_int), ("rows", c_int), ("cols", c_int), ("stride_r", c_int64),
                ("stride_c", c_int64), ("src", c_void_p), ("src2", c_void_p), ("dst", c_void_p), ("dst2", c_void_p),
                ("dst3", c_void_p), ("f0", c_float), ("f1", c_float), ("f2", c_float), ("tap_src", ctypes.c_ubyte * 28)]


PACK_TAPS, PACK_C3_K, PACK_C3_N, PACK_GDN = range(4)
PACK_MAX_JOBS = 48


class DistTerms(Structure):
    _fields_ = [("dist", c_void_p * 4), ("n", c_int * 4), ("scale", c_float * 4), ("n_terms", c_int)]


_SIGNATURES = {
    "hpc_version": (c_int, []),
    "hpc_last_error": (c_char_p, []),
    "hpc_device_is_sm100": (c_int, []),
    "hpc_launch_count": (ctypes.c_ulonglong, []),
    "hpc_prof_enable": (c_int, [c_int]),
    "hpc_prof_pause": (c_int, [c_int]),
    "hpc_prof_read": (c_int, [c_int, POINTER(c_double), POINTER(ctypes.c_ulonglong), POINTER(c_double)]),
    "hpc_prof_records": (c_int, [POINTER(c_int), POINTER(ctypes.c_float), POINTER(c_double), c_int]),
    "hpc_bound_fwd": (c_int, [c_void_p, c_void_p, c_int64, c_float, c_int, c_void_p]),
    "hpc_bound_bwd": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_float, c_int, c_void_p]),
    "hpc_nonneg_fwd": (c_int, [c_void_p, c_void_p, c_int64, c_float, c_float, c_void_p]),
    "hpc_nonneg_bwd": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_float, c_float, c_void_p]),
    "hpc_factorized_fwd": (c_int, [c_void_p, c_void_p, POINTER(CdfParams), c_int64, c_int, c_int, c_float,
                                   c_uint64, c_uint64, c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_void_p,
                                   c_void_p]),
    "hpc_factorized_bwd": (c_int, [c_void_p, POINTER(CdfParams), c_int64, c_int, c_int, c_float, c_void_p,
                                   c_float, c_void_p, c_void_p, c_void_p, POINTER(CdfParams), c_void_p, c_size_t,
                                   c_void_p]),
    "hpc_factorized_bwd_scratch_bytes": (c_size_t, [c_int64, c_int]),
    "hpc_cdf_logits_fwd": (c_int, [c_void_p, POINTER(CdfParams), c_int64, c_int, c_void_p, c_void_p]),
    "hpc_cdf_logits_bwd": (c_int, [c_void_p, POINTER(CdfParams), c_int64, c_int, c_void_p, c_void_p,
                                   POINTER(CdfParams), c_void_p]),
    "hpc_cdf_layer_fwd": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int, c_int, c_void_p, c_void_p]),
    "hpc_cdf_layer_bwd": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int, c_int, c_void_p, c_void_p,
                                  c_void_p, c_void_p, c_void_p, c_void_p]),
    "hpc_conditional_fwd": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int, c_float, c_float,
                                    c_uint64, c_uint64, c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_void_p,
                                    c_void_p]),
    "hpc_conditional_bwd": (c_int, [c_void_p, c_void_p, c_int64, c_int, c_int, c_float, c_float, c_void_p,
                                    c_float, c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p, c_void_p,
                                    c_int, c_void_p]),
    "hpc_exp_clamp_bwd": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int64, c_void_p]),
    "hpc_bits_fwd": (c_int, [c_void_p, c_int64, c_void_p, c_void_p, c_void_p]),
    "hpc_bits_bwd": (c_int, [c_void_p, c_void_p, c_int64, c_void_p, c_void_p]),
    "hpc_reduce_scratch_bytes": (c_size_t, []),
    "hpc_clamp01_sse_fwd": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int64, c_float, c_int, c_void_p,
                                    c_void_p, c_void_p]),
    "hpc_clamp01_bwd": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_float, c_void_p, c_int, c_void_p,
                                c_void_p]),
    "hpc_abs_fwd": (c_int, [c_void_p, c_void_p, c_int64, c_void_p]),
    "hpc_abs_bwd": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_void_p]),
    "hpc_relu_mask": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int64, c_void_p]),
    "hpc_psnr_from_sse": (c_int, [c_void_p, c_int, c_int64, c_float, c_void_p, c_void_p]),
    "hpc_ssim_level_fwd": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_int, c_float,
                                   c_float, c_float, c_void_p, c_void_p, c_void_p, c_void_p]),
    "hpc_ssim_level_bwd": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_int, c_float,
                                   c_float, c_float, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_void_p]),
    "hpc_avgpool2_fwd": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_float, c_void_p]),
    "hpc_avgpool2_bwd": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_float, c_int, c_void_p]),
    "hpc_msssim_combine_fwd": (c_int, [c_void_p, POINTER(c_double), c_int, c_int, POINTER(c_float), c_int,
                                       c_int, c_float, c_void_p, c_void_p]),
    "hpc_msssim_combine_bwd": (c_int, [c_void_p, POINTER(c_double), c_int, c_int, POINTER(c_float), c_int,
                                       c_int, c_float, c_void_p, c_float, c_void_p, c_void_p]),
    "hpc_igemm_fwd_workspace_bytes": (c_size_t, [POINTER(IgemmFwdParams)]),
    "hpc_igemm_fwd": (c_int, [POINTER(IgemmFwdParams), c_void_p]),
    "hpc_igemm_wgrad": (c_int, [POINTER(IgemmWgradParams), c_void_p]),
    "hpc_igemm_wgrad_workspace_bytes": (c_size_t, [POINTER(IgemmWgradParams)]),
    "hpc_igemm_fwd_plan_describe": (c_int, [POINTER(IgemmFwdParams), POINTER(c_int), c_int]),
    "hpc_igemm_wgrad_plan_describe": (c_int, [POINTER(IgemmWgradParams), POINTER(c_int), c_int]),
    "hpc_c3_kmaj_source": (c_int, [c_int]),
    "hpc_pack_weight": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int64, c_int64,
                                POINTER(c_int), c_void_p]),
    "hpc_pack_multi": (c_int, [POINTER(PackJob), c_int, c_int, c_void_p]),
    "hpc_u8_crop_normalize": (c_int, [c_void_p, c_int64, c_int64, c_int, c_int, c_int, c_void_p, c_void_p, c_int, c_int, c_int, c_int,
                                      POINTER(c_float), POINTER(c_float), c_void_p, c_void_p]),
    "hpc_colsum_static": (c_int, [c_void_p, c_int, c_int64, c_int, c_float, c_void_p, c_void_p]),
    "hpc_unpack_wgrad": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int64, c_int64, POINTER(c_int),
                                 c_int, c_void_p]),
    "hpc_colsum": (c_int, [c_void_p, c_int, c_int64, c_int, c_float, c_void_p, c_void_p, c_size_t, c_void_p]),
    "hpc_colsum_scratch_bytes": (c_size_t, [c_int64, c_int]),
    "hpc_square": (c_int, [c_void_p, c_void_p, c_int, c_int64, c_int, c_void_p, c_void_p]),
    "hpc_gdn_bwd_t": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int64, c_int, c_void_p]),
    "hpc_nchw_to_nhwc": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p]),
    "hpc_nhwc_to_nchw": (c_int, [c_void_p, c_int, c_void_p, c_int, c_int, c_int, c_int, c_void_p]),
    "hpc_cast": (c_int, [c_void_p, c_int, c_void_p, c_int, c_int64, c_void_p]),
    "hpc_adamw_step": (c_int, [POINTER(AdamwSeg), c_int, c_double, c_double, c_double, c_double, c_int, c_void_p]),
    "hpc_conv_c3_fwd": (c_int, [c_void_p, c_void_p, c_int64, c_int64, c_void_p, c_int, c_int, c_int, c_int,
                                c_int, c_void_p, c_void_p, c_void_p, c_int, c_void_p]),
    "hpc_deconv_c3_fwd": (c_int, [c_void_p, c_int, c_void_p, c_int64, c_int64, c_void_p, c_int, c_int, c_int,
                                  c_int, c_int, c_void_p, c_void_p]),
    "hpc_wgrad_c3": (c_int, [c_void_p, c_int, c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p, c_int64,
                             c_int64, c_void_p, c_size_t, c_void_p]),
    "hpc_wgrad_c3_scratch_bytes": (c_size_t, [c_int, c_int, c_int, c_int, c_int]),
    "hpc_gdn_fwd_umma": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int, c_void_p, c_void_p, c_void_p]),
    "hpc_gdn_bwd_fused_umma_workspace_bytes": (c_size_t, [c_int64, c_int]),
    "hpc_gdn_bwd_fused_umma": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int, c_void_p,
                                       c_void_p, c_float, c_void_p, c_float, c_void_p, c_void_p, c_void_p,
                                       c_void_p, c_size_t, c_void_p]),
    "hpc_gdn_bwd_fused_rc_umma": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int, c_void_p,
                                       c_void_p, c_float, c_void_p, c_float, c_void_p, c_void_p, c_void_p,
                                       c_void_p, c_size_t, c_void_p]),
    "hpc_gdn_prepare": (c_int, [c_void_p, c_void_p, c_int, c_float, c_float, c_float, c_void_p, c_void_p, c_void_p,
                                c_void_p]),
    "hpc_gdn_bwd_umma_parts": (c_int, [c_int64]),
    "hpc_gdn_bwd_umma": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int, c_void_p, c_void_p,
                                 c_void_p, c_void_p, c_void_p]),
    "hpc_pack_c3_weight": (c_int, [c_void_p, c_int64, c_int64, c_int, c_void_p, c_void_p, c_void_p]),
    "hpc_conv_c3_fwd_umma": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p,
                                     c_void_p]),
    "hpc_deconv_c3_fwd_umma": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p]),
    "hpc_wgrad_c3_umma": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_int64, c_int64, c_void_p,
                                  c_size_t, c_void_p]),
    "hpc_wgrad_c3_umma_workspace_bytes": (c_size_t, [c_int, c_int, c_int, c_int]),
    "hpc_nchw_channel_sum": (c_int, [c_void_p, c_int, c_int, c_int64, c_void_p, c_void_p, c_void_p]),
    "hpc_rans_stream_capacity": (c_int, [c_int, c_int]),
    "hpc_rans_encode": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p,
                                c_void_p, c_void_p]),
    "hpc_rans_compact": (c_int, [c_void_p, c_int, c_void_p, c_void_p, c_int, c_void_p, c_void_p]),
    "hpc_rans_decode": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int,
                                c_void_p, c_void_p, c_int, c_void_p, c_void_p]),
    "hpc_scale_levels": (c_int, [c_void_p, c_int64, c_float, c_float, c_int, c_void_p, c_void_p]),
    "hpc_minmax_rint": (c_int, [c_void_p, c_int64, c_void_p, c_void_p]),
    "hpc_rd_combine": (c_int, [c_void_p, c_void_p, POINTER(DistTerms), c_float, c_float, c_void_p, c_void_p]),
}

EXPORTED_SYMBOLS = tuple(_SIGNATURES)
_lib = None


def lib():
    """The loaded library (ctypes.CDLL) with argtypes set.  Raises if it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: the CUDA extension is the only implementation of this package "
                "(no CPU fallback). Build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "or `python -m image_compression_b200._lib`.")
        l = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in _SIGNATURES.items():
            fn = getattr(l, name)
            fn.restype = res
            fn.argtypes = args
        _lib = l
    return _lib


class HpcError(RuntimeError):
    pass


def check(rc):
    if rc != 0:
        msg = lib().hpc_last_error()
        raise HpcError(f"hpcodec error {rc}: {msg.decode() if msg else '?'}")


_prof_state = {"on": False}


def prof_enable(on):
    """Start (clearing earlier records) or stop the kernel-class event timing."""
    check(lib().hpc_prof_enable(int(on)))
    _prof_state["on"] = bool(on)


def prof_is_on():
    return _prof_state["on"]


def prof_pause(paused):
    """Suspend / resume recording without clearing the records (used around CUDA-graph capture)."""
    check(lib().hpc_prof_pause(int(paused)))
    _prof_state["on"] = not paused


def prof_read(cls):
    """(ms_total, launches, work_total) of one kernel class since prof_enable(True)."""
    ms, n, w = c_double(), ctypes.c_ulonglong(), c_double()
    check(lib().hpc_prof_read(cls, byref(ms), byref(n), byref(w)))
    return ms.value, n.value, w.value


def prof_records():
    """[(class name, ms, work)] of every record since prof_enable(True), in call order."""
    n = lib().hpc_prof_records(None, None, None, 0)
    cls, ms, work = (c_int * n)(), (ctypes.c_float * n)(), (c_double * n)()
    n = min(n, lib().hpc_prof_records(cls, ms, work, n))
    return [(PROF_NAMES[cls[i]], float(ms[i]), float(work[i])) for i in range(n)]


def ptr(t):
    """Raw device pointer of a tensor (None -> NULL)."""
    return None if t is None else t.data_ptr()


def stream():
    """Raw handle of torch's current stream on the current device (torch.cuda.current_stream() builds a Stream
    object on every call: ~10x the cost, and there are ~200 launches per training step)."""
    return torch._C._cuda_getCurrentRawStream(torch.cuda.current_device())


def dtype_code(t_or_dtype):
    dt = t_or_dtype.dtype if isinstance(t_or_dtype, torch.Tensor) else t_or_dtype
    if dt == torch.float32:
        return F32
    if dt == torch.bfloat16:
        return BF16
    raise TypeError(f"unsupported dtype {dt}")


def require_cuda(*tensors):
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise RuntimeError(
                "image_compression_b200 runs only on CUDA tensors (sm_100a); got a CPU tensor. "
                "There is no CPU fallback by design.")


if __name__ == "__main__":
    print(build_library(force="--force" in sys.argv, verbose=True))
