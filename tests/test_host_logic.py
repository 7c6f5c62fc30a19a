"""CPU tests of the host-side plans: tap tables, phase decomposition, weight packing."""
import pytest
import torch
import torch.nn.functional as F

import igemm_emulator as E
from image_compression_b200 import _lib
from image_compression_b200 import hotpath as hp

torch.manual_seed(0)
CASES = [(False, 5, 2), (False, 3, 1), (True, 5, 2), (True, 3, 1)]


def _ref(transposed, x, w, k, s):
    if transposed:
        return F.conv_transpose2d(x, w, None, stride=s, padding=k // 2, output_padding=s - 1)
    return F.conv2d(x, w, None, stride=s, padding=k // 2)


@pytest.mark.parametrize("transposed,k,s", CASES)
def test_forward_dgrad_wgrad_plans(transposed, k, s):
    ci, co, N, H, W = 3, 4, 2, 6, 8
    x = torch.randn(N, ci, H, W, dtype=torch.float64, requires_grad=True)
    w = torch.randn((ci, co, k, k) if transposed else (co, ci, k, k), dtype=torch.float64, requires_grad=True)
    y = _ref(transposed, x, w, k, s)
    g = torch.randn_like(y)
    gx_ref, gw_ref = torch.autograd.grad(y, (x, w), g)
    spec = hp.ConvSpec(transposed, ci, co, k, s)
    K = k * k
    OH, OW = spec.out_hw(H, W)
    assert (OH, OW) == tuple(y.shape[2:])
    xn, gn = x.detach().permute(0, 2, 3, 1), g.permute(0, 2, 3, 1)
    wd = w.detach()
    # forward (ConvSpec.forward)
    if transposed:
        taps, phases = spec.deconv_taps
        wp = E.pack_weight(wd, taps, co, ci, K, co * K)
        out = E.igemm_fwd(xn, wp, taps, phases, (H, W), 1, s, (OH, OW), co)
    else:
        taps, phases = spec.conv_taps
        wp = E.pack_weight(wd, taps, co, ci, ci * K, K)
        out = E.igemm_fwd(xn, wp, taps, phases, (OH, OW), s, 1, (OH, OW), co)
    assert torch.allclose(out.permute(0, 3, 1, 2), y.detach(), atol=1e-10)
    # dgrad (ConvSpec.dgrad)
    if transposed:
        taps, phases = spec.conv_taps
        wp = E.pack_weight(wd, taps, ci, co, co * K, K)
        gx = E.igemm_fwd(gn, wp, taps, phases, (H, W), s, 1, (H, W), ci)
    else:
        taps, phases = spec.deconv_taps
        wp = E.pack_weight(wd, taps, ci, co, K, ci * K)
        gx = E.igemm_fwd(gn, wp, taps, phases, (OH, OW), 1, s, (H, W), ci)
    assert torch.allclose(gx.permute(0, 3, 1, 2), gx_ref, atol=1e-10)
    # wgrad (ConvSpec.wgrad)
    taps, phases = spec.deconv_taps
    if transposed:
        gw = E.igemm_wgrad(xn, gn, taps, phases, s)
        gw = E.unpack_wgrad(gw, w.shape, taps, ci, co, co * K, K)
    else:
        gw = E.igemm_wgrad(gn, xn, taps, phases, s)
        gw = E.unpack_wgrad(gw, w.shape, taps, co, ci, ci * K, K)
    assert torch.allclose(gw, gw_ref, atol=1e-10)


def test_phase_tap_counts():
    taps, phases = hp.deconv_type_taps(5, 2, 2)
    assert [p[0] for p in phases] == [9, 6, 6, 4] and len(taps) == 25
    taps, phases = hp.conv_type_taps(5, 2, 2)
    assert len(taps) == 25 and phases == [(25, 0, 0, 0)]


def test_gdn_bounds_match_reference_quirk():
    assert hp.PEDESTAL == 2.0 ** -36
    assert hp.GAMMA_BOUND == 2.0 ** -36          # sqrt(0 + pedestal^2): gdn.py:53
    assert abs(hp.beta_bound(1e-6) - 1e-3) < 1e-6


def test_gauss_filter_is_the_separable_factor():
    import sys
    from oracle import hyperprior_oracle as O
    g2 = O.gauss_filter_2d(11, 1.5)[0, 0]
    g1 = hp.gauss_filter_1d(11, 1.5, "cpu")
    assert torch.allclose(torch.outer(g1, g1), g2, atol=1e-8)


def test_cpu_tensor_is_rejected_loudly():
    import image_compression_b200 as pkg
    m = pkg.build_model(pkg.get_recipe_cfg("psnr_256"))
    with pytest.raises(RuntimeError, match="CUDA"):
        m(torch.rand(1, 3, 64, 64))


def test_fused_adamw_host_logic():
    """solver.make_optimizer mirrors solver/optim.py:20-46 (one group per tensor, the "bias" rule) and refuses CPU
    tensors (no fallback); non-adamw optimizers are left to the reference."""
    import image_compression_b200 as pkg
    from image_compression_b200.solver import FusedAdamW, make_optimizer
    cfg = pkg.get_recipe_cfg("psnr_256")
    model = pkg.build_model(cfg)
    opt = make_optimizer(cfg, model)
    assert isinstance(opt, FusedAdamW) and len(opt.param_groups) == 50
    names = [n for n, _ in model.named_parameters()]
    for n, g in zip(names, opt.param_groups):
        assert g["weight_decay"] == (cfg.SOLVER.WEIGHT_DECAY_BIAS if "bias" in n else cfg.SOLVER.WEIGHT_DECAY)
        assert g["lr"] == cfg.SOLVER.BASE_LR and g["eps"] == cfg.SOLVER.EPS
    assert sum("bias" in n for n in names) == 17
    for p in model.parameters():
        p.grad = torch.zeros_like(p)
    with pytest.raises(RuntimeError, match="CUDA"):
        opt.step()
    cfg.SOLVER.OPT_NAME = "sgd"
    with pytest.raises(NotImplementedError):
        make_optimizer(cfg, model)


def test_early_all_reduce_slot_range(monkeypatch):
    """FlatGradients._reduce_early (the hook the backward schedules call after every layer): slots become reducible in
    buffer order -- the range between the frontier and the first slot not yet reported is all-reduced once it holds
    >= min_bucket elements or at a flush; a gradient that is not a slot view is left to the final all-reduce, which
    covers everything beyond the frontier."""
    import torch.distributed as dist
    from image_compression_b200 import sharding
    params = [torch.nn.Parameter(torch.zeros(n)) for n in (5, 16, 7, 9, 3)]
    fg = sharding.FlatGradients(params, torch.device("cpu"), overlap=False)
    fg.min_bucket = 20
    calls = []

    class _Work:
        def wait(self):
            return True

    def fake_all_reduce(t, op=None, group=None, async_op=False):
        calls.append(((t.data_ptr() - fg.flat.data_ptr()) // 4, t.numel()))
        return _Work()
    monkeypatch.setattr(dist, "all_reduce", fake_all_reduce)
    monkeypatch.setattr(dist, "is_initialized", lambda: True)
    monkeypatch.setattr(dist, "get_world_size", lambda group=None: 2)
    monkeypatch.setattr(dist, "get_backend", lambda group=None: "nccl")
    views = [fg._view(p) for p in params]
    fg._reduce_early([views[2], views[3]])                     # slots 2, 3 ready, but slot 0 is the frontier: nothing yet
    assert calls == [] and fg._frontier == 0
    fg._reduce_early([views[0]])                               # slot 0 alone: 8 padded elements < min_bucket
    assert calls == [] and fg._frontier == 0
    fg._reduce_early([views[1]])                               # 0..3 contiguous now: one all-reduce up to slot 4's offset
    assert calls == [(0, fg.offsets[4])] and fg._frontier == 4
    fg._reduce_early([torch.zeros(3)])                         # a stray gradient (not a view of the flat buffer): ignored
    assert len(calls) == 1 and fg._frontier == 4
    for p, v in zip(params, views):
        p.grad = v
    fg.all_reduce_mean()                                       # the rest: slot 4 to the end of the buffer; state reset
    assert calls[1] == (fg.offsets[4], fg.flat.numel() - fg.offsets[4]) and fg._frontier == 0 and not fg._early
    calls.clear()
    fg._reduce_early([views[0]], flush=True)                   # a flush starts the reduce of whatever is contiguous
    assert calls == [(0, fg.offsets[1])] and fg._frontier == 1

@pytest.mark.parametrize("transposed,k,s,grid", [(False, 5, 2, (64, 64)), (True, 5, 2, (64, 64)), (False, 3, 1, (16, 16)),
                                                 (False, 5, 2, (4, 4)), (True, 5, 2, (12, 12))])
def test_wgrad_unit_plan_covers_every_block_once_and_shares_boxes(transposed, k, s, grid):
    """Host logic of the tcgen05 weight-gradient engine (igemm_umma.cu::wgrad_plan, no GPU needed): every (tap, r-chunk)
    block of the result is owned by exactly one unit; an atom lies inside one of its unit's `lo` boxes at the row that
    matches its dh; boxes do not overlap in the stage; on 8x8 K blocks the taps of one (dw, r-chunk) share a box
    (18 KB of `lo` per full unit instead of 32 KB), on smaller grids (several images per K block) they do not."""
    import ctypes
    spec = hp.ConvSpec(transposed, 192, 192, k, s)
    taps, phases = spec.deconv_taps
    GH, GW = grid
    P = _lib.IgemmWgradParams()
    P.engine, P.dtype = _lib.ENGINE_UMMA, _lib.BF16
    P.lo, P.hi, P.gw_f32 = 1 << 20, 2 << 20, 3 << 20                  # never dereferenced by the planner
    P.N, P.GH, P.GW, P.R = 64, GH, GW, 192
    P.HH, P.HW, P.Cc, P.hs = GH * s, GW * s, 192, s
    hp._fill_taps(P, taps, phases)
    P.scale = 1.0
    buf = (ctypes.c_int * (4 + 35 * 32))()
    n = _lib.lib().hpc_igemm_wgrad_plan_describe(ctypes.byref(P), buf, len(buf))
    assert n > 0, _lib.lib().hpc_last_error()
    nunits, nsplits, a_bytes, kblocks = buf[0], buf[1], buf[2], buf[3]
    assert n == 4 + 35 * nunits and 1 <= nsplits and nunits * nsplits <= 148 and kblocks >= nsplits
    share = GH >= 8 and GW >= 8
    th = 8 if share else min(GH, 8)
    seen, per_unit_kb = set(), []
    for u in range(nunits):
        r = list(buf[4 + 35 * u: 4 + 35 * (u + 1)])
        ph, natoms, nloads = r[:3]
        atoms = [tuple(r[3 + 3 * a: 6 + 3 * a]) for a in range(4)]
        loads = [tuple(r[15 + 5 * l: 20 + 5 * l]) for l in range(4)]
        assert 1 <= natoms <= 4 and 1 <= nloads <= natoms
        assert all(x == (-1, -1, -1) for x in atoms[natoms:]) and all(x == (-1,) * 5 for x in loads[nloads:])
        nt, begin = phases[ph][0], phases[ph][1]
        ends = []
        for rc, dw, dh0, rows, off in loads[:nloads]:
            assert th <= rows <= th + 2
            assert all(off >= e for e in ends)                         # boxes in ascending, non-overlapping order
            ends.append(off + (rows if share else 8))                  # KB: one 8-pixel row per KB | a whole 64-pixel box
        offs = []
        for tap, rc, off in atoms[:natoms]:
            assert begin <= tap < begin + nt and (tap, rc) not in seen
            seen.add((tap, rc))
            dh, dw = taps[tap][0], taps[tap][1]
            home = [(lrc, ldw, ldh, rows, loff) for lrc, ldw, ldh, rows, loff in loads[:nloads]
                    if lrc == rc and ldw == dw and ldh <= dh < ldh + (rows - th) + 1 and off == loff + (dh - ldh) * (1 if share else 0)]
            assert len(home) == 1, (u, tap, rc, off, loads)
            offs.append(off)
        assert offs == sorted(offs)                                    # the second atom of an MMA lies at a higher address
        per_unit_kb.append(max(ends))
        assert max(ends) * 1024 <= a_bytes
    assert seen == {(t, rc) for t in range(len(taps)) for rc in range(3)}
    full = [kb for u, kb in enumerate(per_unit_kb) if buf[4 + 35 * u + 1] == 4]
    assert max(full) == (18 if share else 32), (share, full)


def test_c3_kmaj_axis_is_a_zero_padded_permutation():
    """k' = (cs*5 + kh)*8 + kw of the packed 3-channel conv weights (thin_umma.cu): the 75 live positions map onto the
    reference's (cs, kh*5+kw) axis one to one, the other 53 hold zeros."""
    src = [_lib.lib().hpc_c3_kmaj_source(k) for k in range(128)]
    live = [v for v in src if v >= 0]
    assert sorted(live) == list(range(75)) and src.count(-1) == 53
    for cs in range(3):
        for kh in range(5):
            for kw in range(8):
                want = cs * 25 + kh * 5 + kw if kw < 5 else -1
                assert src[(cs * 5 + kh) * 8 + kw] == want
    assert all(v == -1 for v in src[120:])


def test_data_parallel_replicas_are_refused_loudly():
    """nn.DataParallel over several GPUs replicates the module onto threads of one process (torch marks replicas with
    `_is_replica`); the hot path's caches are per process, so the model refuses instead of racing."""
    import image_compression_b200 as pkg
    model = pkg.build_model(pkg.get_recipe_cfg("psnr_256"))
    model._is_replica = True
    with pytest.raises(RuntimeError, match="one process per GPU"):
        model(torch.zeros(1, 3, 64, 64))


def _fwd_plan(transposed, k, s, grid, N=64, C=192, dgrad=False):
    import ctypes
    spec = hp.ConvSpec(transposed, C, C, k, s)
    GH, GW = grid
    # forward of a conv / dgrad of a transposed conv: strided input, one phase; forward of a transposed conv / dgrad of a
    # conv: unit input stride, s*s output phases (hotpath.ConvSpec.forward / dgrad)
    conv_like = (not transposed) != dgrad
    taps, phases = spec.conv_taps if conv_like else spec.deconv_taps
    is_, os_ = (s, 1) if conv_like else (1, s)
    P = _lib.IgemmFwdParams()
    P.engine, P.dtype = _lib.ENGINE_UMMA, _lib.BF16
    P.x, P.w, P.out_lp = 1 << 20, 2 << 20, 3 << 20                    # never dereferenced by the planner
    P.N, P.H, P.W, P.Ci = N, GH * is_, GW * is_, C
    P.GH, P.GW, P.is_, P.os = GH, GW, is_, os_
    P.OH, P.OW, P.Co = GH * os_, GW * os_, C
    hp._fill_taps(P, taps, phases)
    P.epi = _lib.EPI_LINEAR
    wsb = _lib.lib().hpc_igemm_fwd_workspace_bytes(ctypes.byref(P))
    if wsb:
        P.workspace, P.workspace_bytes = 4 << 20, wsb
    buf = (ctypes.c_int * 12)()
    assert _lib.lib().hpc_igemm_fwd_plan_describe(ctypes.byref(P), buf, 12) == 12, _lib.lib().hpc_last_error()
    keys = ("pair", "mt", "share", "ksplit", "interleave", "stage_out", "tw", "th", "nb", "tiles_per_phase", "tiles", "grid")
    return dict(zip(keys, list(buf)))


def test_forward_engine_plan_for_the_model_layers_at_the_benchmark_size():
    """Which build of the tcgen05 forward engine each 192 -> 192 layer of a 64 x 256x256 step gets (igemm_umma.cu::fwd_plan,
    no GPU needed): DESIGN 3.1's table.  grid = the layer's iteration grid (output pixels of a conv, input pixels of a
    transposed conv)."""
    big = dict(pair=1, mt=2, ksplit=1, stage_out=1, tw=16, th=8, nb=1)
    for grid, tiles in (((64, 64), 2048), ((32, 32), 512)):            # g_a conv2, conv3: groups of four tiles per CTA pair
        pl = _fwd_plan(False, 5, 2, grid)
        assert {k: pl[k] for k in big} == big and pl["tiles"] == tiles and pl["grid"] == 148 and pl["interleave"] == 0
    pl = _fwd_plan(False, 5, 2, (16, 16))                              # g_a conv4: 128 tiles, one per SM, no pairs
    assert (pl["pair"], pl["mt"], pl["ksplit"], pl["tiles"], pl["grid"]) == (0, 1, 1, 128, 128)
    pl = _fwd_plan(False, 3, 1, (16, 16))                              # h_a conv1 (3x3, stride 1)
    assert (pl["pair"], pl["tiles"], pl["grid"]) == (0, 128, 128)
    pl = _fwd_plan(False, 5, 2, (8, 8))                                # h_a conv2: 32 tiles of two images, K split 4 ways
    assert (pl["pair"], pl["nb"], pl["tiles"], pl["ksplit"], pl["grid"]) == (0, 2, 32, 4, 128)
    pl = _fwd_plan(False, 5, 2, (4, 4))                                # h_a conv3: 8 tiles of eight images, K split 16 ways
    assert (pl["pair"], pl["tw"], pl["th"], pl["nb"], pl["tiles"], pl["ksplit"], pl["grid"]) == (0, 4, 4, 8, 8, 16, 128)
    # transposed convs: four output phases over the INPUT grid
    pl = _fwd_plan(True, 5, 2, (16, 16))                               # g_s deconv1: 4 x 128 tiles -> pairs, 32 groups per phase
    assert (pl["pair"], pl["mt"], pl["tiles_per_phase"], pl["tiles"], pl["interleave"]) == (1, 2, 128, 512, 0)
    pl = _fwd_plan(True, 5, 2, (32, 32))                               # deconv2: 1.73 rounds per phase: phase-major
    assert (pl["pair"], pl["mt"], pl["tiles_per_phase"], pl["interleave"], pl["grid"]) == (1, 2, 512, 0, 148)
    pl = _fwd_plan(True, 5, 2, (64, 64))                               # deconv3: 6.9 rounds: phases interleaved per tile group
    assert (pl["pair"], pl["mt"], pl["tiles_per_phase"], pl["interleave"]) == (1, 2, 2048, 1)
    # dgrads mirror them: dgrad of a strided conv is phase-structured, dgrad of a transposed conv is conv-like
    assert _fwd_plan(False, 5, 2, (64, 64), dgrad=True)["interleave"] == 1
    assert _fwd_plan(True, 5, 2, (64, 64), dgrad=True) == _fwd_plan(False, 5, 2, (64, 64))
    assert all(_fwd_plan(t, 5, 2, g)["share"] == 0 for t in (False, True) for g in ((64, 64), (4, 4)))   # off by default
