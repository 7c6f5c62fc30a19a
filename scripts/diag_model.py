import sys
import torch
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import image_compression_b200 as pkg
from oracle import hyperprior_oracle as O
from util import make_pair, perturb

H, W, N = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
model, params, ocfg = make_pair("psnr_256", 192)
params = perturb(params); model.load_state_dict(params); model.train()
torch.manual_seed(21)
x = torch.rand(N, 3, H, W)
nz = torch.rand(N, 192, H // 64, W // 64); ny = torch.rand(N, 192, H // 16, W // 16)
with torch.no_grad():
    _, lo = O.compressor2018_forward(params, x, ocfg, True, nz, ny)
print("oracle", {k: round(float(v), 6) for k, v in lo.items()})
for mode in ("fp32", "bf16", "bf16_simt"):
    for rep in range(2):
        with pkg.compute_mode(mode), torch.no_grad():
            _, l = model(x.cuda(), nz.cuda(), ny.cuda())
        print(mode, rep, {k: round(float(v), 6) for k, v in l.items()})
