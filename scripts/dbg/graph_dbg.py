import torch, image_compression_b200 as pkg
from image_compression_b200 import utils as U
from image_compression_b200.utils.graphs import GraphedCallable
torch.manual_seed(3)
model = pkg.build_model(pkg.get_recipe_cfg("psnr_256")).to("cuda").eval()
x = torch.rand(1, 3, 256, 256, device="cuda")
def step(t):
    xt, losses = model(t)
    ps, ms = U.image_metrics(xt, t)
    return xt, losses["bpp"], losses["z_entropy"], losses["y_entropy"], losses["total_loss"], ps, ms
with torch.no_grad():
    a = [v.clone() for v in step(x)]
    b = [v.clone() for v in step(x)]
    print("eager==eager", [bool(torch.equal(u, v)) for u, v in zip(a, b)])
    g = GraphedCallable(step, x)
    for i in range(3):
        c = g(x)
        print("graph==eager", [bool(torch.equal(u, v)) for u, v in zip(c, a)], [float((u.float()-v.float()).abs().max()) for u, v in zip(c, a)])
