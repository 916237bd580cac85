"""Developer tool: a few forward+backward passes of the per-frame SMPL table kernels (for `ncu`)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from moss_b200 import skinning as K, synth
dev = torch.device("cuda:0")
body = synth.body()
V = body.verts.shape[0]
smpl = K.Smpl(body.verts.to(dev), body.shapedirs.to(dev), body.posedirs.t().reshape(V, 3, 207).contiguous().to(dev),
              body.J_regressor.to(dev), body.weights.to(dev), body.parents.to(dev))
fr = synth.frame_params(7)
poses = fr.params["poses"].to(dev).requires_grad_(True)
shapes = fr.params["shapes"].to(dev)
gA, gpo = torch.randn(24, 4, 4, device=dev), torch.randn(V, 3, device=dev)
flush = torch.empty(256 * 1024 * 1024 // 4, dtype=torch.float32, device=dev)
for i in range(int(sys.argv[1]) if len(sys.argv) > 1 else 6):
    if i % 2 == 0:
        flush.zero_()              # even passes start with a cold L2
    A, po, so = K.smpl_frame(smpl, poses, shapes)
    torch.autograd.backward([A, po], [gA, gpo])
torch.cuda.synchronize()
print("ok")
# the raw entry points without the posedirs pass: forward = shape blocks + chain only, backward = ONE block (chain only)
import ctypes as C
from moss_b200 import _abi
lib = _abi.load()
scratch = torch.zeros(lib.moss_smpl_frame_scratch_bytes(V), dtype=torch.uint8, device=dev)
par = smpl.parents.to(torch.int32)
p72 = poses.detach().reshape(-1).contiguous()
f = K._fill_frame(smpl, p72, shapes.reshape(-1), None, scratch, par)
A, J, dp = torch.empty(24, 16, device=dev), torch.empty(24, 3, device=dev), torch.empty(72, device=dev)
for i in range(3):
    flush.zero_()
    lib.moss_smpl_frame_forward(None, C.byref(f), A.data_ptr(), None, J.data_ptr(), None, None)
    lib.moss_smpl_frame_backward(None, C.byref(f), A.data_ptr(), J.data_ptr(), gA.data_ptr(), None, dp.data_ptr(), None)
torch.cuda.synchronize()
print("ok raw")
