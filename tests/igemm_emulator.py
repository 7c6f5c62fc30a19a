"""Pure-torch CPU emulation of the implicit-GEMM contracts in include/hpcodec.h
(hpc_igemm_fwd / hpc_igemm_wgrad / hpc_pack_weight / hpc_unpack_wgrad).  Used by the CPU tests to
prove that the tap tables and weight packings built in image_compression_b200/hotpath.py reproduce
Conv2d / ConvTranspose2d and their gradients before any kernel runs."""
import torch


def pack_weight(w, taps, rows, cols, stride_r, stride_c):
    flat = w.reshape(-1)
    out = torch.empty(len(taps), rows, cols, dtype=w.dtype)
    r = torch.arange(rows).view(-1, 1)
    c = torch.arange(cols).view(1, -1)
    for t, (_, _, src) in enumerate(taps):
        out[t] = flat[r * stride_r + c * stride_c + src]
    return out


def unpack_wgrad(gw, shape, taps, rows, cols, stride_r, stride_c):
    out = torch.zeros(shape, dtype=gw.dtype).reshape(-1)
    r = torch.arange(rows).view(-1, 1)
    c = torch.arange(cols).view(1, -1)
    for t, (_, _, src) in enumerate(taps):
        out[(r * stride_r + c * stride_c + src).reshape(-1)] = gw[t].reshape(-1)
    return out.view(shape)


def igemm_fwd(x, wp, taps, phases, grid_hw, is_, os_, out_hw, Co):
    """x (N,H,W,Ci) -> (N,OH,OW,Co)"""
    N, H, W, Ci = x.shape
    GH, GW = grid_hw
    OH, OW = out_hw
    out = torch.zeros(N, OH, OW, Co, dtype=x.dtype)
    for (nt, begin, a, b) in phases:
        for i in range(GH):
            oh = i * os_ + a
            if oh >= OH:
                continue
            for j in range(GW):
                ow = j * os_ + b
                if ow >= OW:
                    continue
                acc = torch.zeros(N, Co, dtype=x.dtype)
                for t in range(begin, begin + nt):
                    ih, iw = i * is_ + taps[t][0], j * is_ + taps[t][1]
                    if 0 <= ih < H and 0 <= iw < W:
                        acc += x[:, ih, iw, :] @ wp[t].T
                out[:, oh, ow, :] = acc
    return out


def igemm_wgrad(lo, hi, taps, phases, hs):
    N, GH, GW, R = lo.shape
    _, HH, HW, Cc = hi.shape
    gw = torch.zeros(len(taps), R, Cc, dtype=lo.dtype)
    for (nt, begin, a, b) in phases:
        for t in range(begin, begin + nt):
            dh, dw = taps[t][0], taps[t][1]
            for i in range(GH):
                li, hh = i + dh, i * hs + a
                if not (0 <= li < GH) or hh >= HH:
                    continue
                for j in range(GW):
                    lj, hw = j + dw, j * hs + b
                    if not (0 <= lj < GW) or hw >= HW:
                        continue
                    gw[t] += lo[:, li, lj, :].T @ hi[:, hh, hw, :]
    return gw
