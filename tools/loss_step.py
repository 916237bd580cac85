"""Developer tool: a few passes of the fused image-loss kernel at ZJU size (for `ncu`)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from moss_b200 import losses as ML
from tests.golden.make_loss_golden import synth_case
dev = torch.device("cuda:0")
W, H = (int(v) for v in (sys.argv[1] if len(sys.argv) > 1 else "512x512").split("x"))
box = (W // 4 + 1, H // 9, W - W // 4 - 2, H - H // 7 - 1)
image, alpha, gt, bkgd, bound = (torch.from_numpy(a).to(dev) for a in synth_case(W, H, 11, box))
bounds = ML.mask_bounds(bound)
for _ in range(int(sys.argv[2]) if len(sys.argv) > 2 else 3):
    ig, ag = image.detach().requires_grad_(True), alpha.detach().requires_grad_(True)
    loss, _ = ML.image_losses(ig, ag, gt, bkgd, bound, bounds)
    loss.backward()
torch.cuda.synchronize()
print("ok", float(loss))
