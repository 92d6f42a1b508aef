"""Per-launch milliseconds of one forward (CUDA-event pair around every launch, mean of 5):
python tools/launch_profile.py [model] [batch] [alexnet_geometry]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import dmlc_b200 as b2  # noqa: E402
from oracle import synth  # noqa: E402

model_name = sys.argv[1] if len(sys.argv) > 1 else "resnet18"
B = int(sys.argv[2]) if len(sys.argv) > 2 else 256
geom = int(sys.argv[3]) if len(sys.argv) > 3 else 0
ot = synth.ensure_ot(model_name, os.path.join(ROOT, "tests", "_cache", model_name + ".ot"))
ctx = b2.Context((0,))
model = b2.Model(ctx, model_name, ot, max_batch=B, alexnet_geometry=geom)
x = torch.from_numpy(synth.synthetic_images(B, seed=0)).cuda()
idx = torch.empty((B, 5), dtype=torch.int32, device="cuda")
prob = torch.empty((B, 5), dtype=torch.float32, device="cuda")
st = torch.cuda.current_stream().cuda_stream
acc = {}
for r in range(6):
    rows = model.profile(x.data_ptr(), B, 5, idx.data_ptr(), prob.data_ptr(), st)
    if r:
        for j, (name, ms) in enumerate(rows):
            acc.setdefault((j, name), []).append(ms)
total = 0.0
for (j, name), v in sorted(acc.items()):
    ms = sum(v) / len(v)
    total += ms
    print("%-40s %.4f ms" % (name, ms))
print("sum %.4f ms -> %.0f img/s; %.3f GFLOP/img -> %.0f TFLOP/s" % (total, B / total * 1e3, model.flops_per_image / 1e9,
                                                                  model.flops_per_image * B / total / 1e9))
