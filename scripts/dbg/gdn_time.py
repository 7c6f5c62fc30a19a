"""time the fused GDN kernels alone at the three training shapes"""
import sys, torch
import image_compression_b200 as pkg
from image_compression_b200 import hotpath as hp
C = 192
pkg.set_compute_mode("bf16")
for (N, H, W) in ((64, 128, 128), (64, 64, 64), (64, 32, 32)):
    x = torch.randn(N, H, W, C, device="cuda").to(torch.bfloat16)
    g = torch.randn(N, H, W, C, device="cuda").to(torch.bfloat16)
    gamma = (0.1 * torch.eye(C, device="cuda") + 0.01 * torch.rand(C, C, device="cuda")).view(C, C, 1, 1)
    beta = 1 + torch.rand(C, device="cuda")
    y, nrm = hp.gdn_forward(x, None, gamma, beta, False)
    def t(fn, n=10):
        for _ in range(3): fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n): fn()
        e1.record(); torch.cuda.synchronize()
        return e0.elapsed_time(e1) / n * 1e3
    P = N * H * W
    tf = t(lambda: hp.gdn_forward(x, None, gamma, beta, False))
    tb = t(lambda: hp.gdn_backward(g, x, None, nrm, gamma, False, want_dxsum=True))
    print(f"P={P}: fwd {tf:.1f} us ({3*P*C*2/tf/1e6:.2f} TB/s)  bwd(all launches) {tb:.1f} us ({4*P*C*2/tb/1e6:.2f} TB/s)")
