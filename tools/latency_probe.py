import os, sys, time, numpy as np
sys.path.insert(0, '/root/repo')
import dmlc_b200 as b2
from oracle import synth
ctx = b2.Context((0,))
ot = synth.ensure_ot("resnet18", "/root/repo/tests/_cache/resnet18.ot")
m = b2.Model(ctx, "resnet18", ot, max_batch=256)
pin = b2.PinnedBuffer(256 * b2.IMAGE_BYTES); arr = pin.array.reshape(256,224,224,3); arr[:] = synth.synthetic_images(256, seed=1)
for bs in (1, 8, 64, 256):
    for _ in range(10): m.predict_u8(arr[:bs], k=5)
    ts=[]
    for _ in range(200 if bs<256 else 50):
        t0=time.perf_counter(); m.predict_u8(arr[:bs], k=5); ts.append((time.perf_counter()-t0)*1e3)
    ts.sort(); print(os.environ.get("B2I_NO_PDL","0"), "batch", bs, "p50 %.3f ms p99 %.3f ms" % (ts[len(ts)//2], ts[int(len(ts)*0.99)]))
