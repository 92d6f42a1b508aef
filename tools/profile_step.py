"""One-process driver for ncu: W warm-up steps then S steps of the ResNet-18 bs-256 forward
(device-resident input), nothing else.  Usage: python tools/profile_step.py [W] [S] [tile_n_max] [model]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import dmlc_b200 as b2  # noqa: E402
from oracle import synth  # noqa: E402

W = int(sys.argv[1]) if len(sys.argv) > 1 else 3
S = int(sys.argv[2]) if len(sys.argv) > 2 else 1
TN = int(sys.argv[3]) if len(sys.argv) > 3 else 0
MODEL = sys.argv[4] if len(sys.argv) > 4 else "resnet18"
B = 256
ot = synth.ensure_ot(MODEL, os.path.join(ROOT, "tests", "_cache", MODEL + ".ot"))
ctx = b2.Context((0,))
model = b2.Model(ctx, MODEL, ot, max_batch=B, tile_n_max=TN)
x = torch.from_numpy(synth.synthetic_images(B, seed=0)).cuda()
idx = torch.empty((B, 5), dtype=torch.int32, device="cuda")
prob = torch.empty((B, 5), dtype=torch.float32, device="cuda")
st = torch.cuda.current_stream().cuda_stream
for _ in range(W + S):
    model.forward_device(x.data_ptr(), B, 5, idx.data_ptr(), prob.data_ptr(), None, st)
torch.cuda.synchronize()
print("done", model.launches_per_microbatch)
