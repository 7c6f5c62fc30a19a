"""time the big wgrad / fwd conv kernels alone"""
import torch
import image_compression_b200 as pkg
from image_compression_b200 import hotpath as hp
pkg.set_compute_mode("bf16")
C = 192
def t(fn, n=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3
for (N, H, W, k, s, tr) in ((64, 128, 128, 5, 2, False), (64, 64, 64, 5, 2, False), (64, 32, 32, 5, 2, False), (64, 16, 16, 3, 1, False), (64, 16, 16, 5, 2, False), (64, 8, 8, 5, 2, False), (64, 64, 64, 5, 2, True), (64, 32, 32, 5, 2, True), (64, 16, 16, 5, 2, True)):
    spec = hp.ConvSpec(tr, C, C, k, s)
    OH, OW = spec.out_hw(H, W)
    x = torch.randn(N, H, W, C, device="cuda").to(torch.bfloat16)
    g = torch.randn(N, OH, OW, C, device="cuda").to(torch.bfloat16)
    w = torch.randn(C, C, k, k, device="cuda") * 0.01
    b = torch.zeros(C, device="cuda")
    fl = 2.0 * N * (OH * OW if not tr else H * W) * k * k * C * C
    tw = t(lambda: spec.wgrad(x, g, w))
    tf = t(lambda: spec.forward(x, w, b))
    td = t(lambda: spec.dgrad(g, w, (H, W)))
    print(f"{'deconv' if tr else 'conv'} {H}x{W} k{k}s{s}: wgrad {tw:.1f} us ({fl/tw/1e6:.0f} TF/s)  fwd {tf:.1f} us ({fl/tf/1e6:.0f} TF/s)  dgrad {td:.1f} us ({fl/td/1e6:.0f} TF/s)")
