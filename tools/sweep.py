"""Batch-size sweep (BASELINE.json configs[4]) and the dual-job fair share (configs[3]) on the GPUs
visible to this process.  Writes JSON lines to the file given as argv[1] (default stdout).

  sweep : for each model, batch size 1..4096 and GPU count (1 and all visible), blocking b2i_predict_u8 latency
          (p50 / p99 over the repetitions, host buffers pinned, preprocessing = normalise on the GPU) and images/s;
          plus, per model, b2i_predict_raw_u8 on 500x375 decoded images (resize + crop + normalise on the GPU)
  dual  : ResNet-18 (bs 256) and AlexNet (bs 256) at the same time on disjoint halves of the GPUs
          (reference src/services.rs:199-211) against each job alone on the same half
"""
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import dmlc_b200 as b2  # noqa: E402
from oracle import synth  # noqa: E402

out = open(sys.argv[1], "w") if len(sys.argv) > 1 else sys.stdout
n_gpus = b2.device_count()


def emit(d):
    out.write(json.dumps(d) + "\n")
    out.flush()


def pinned_images(n, seed):
    pin = b2.PinnedBuffer(n * b2.IMAGE_BYTES)
    arr = pin.array.reshape(n, 224, 224, 3)
    arr[:] = synth.synthetic_images(min(n, 256), seed=seed)[np.arange(n) % min(n, 256)]
    return pin, arr


def timed(model, imgs, reps, devices=None):
    for _ in range(3):
        model.predict_u8(imgs, k=5, devices=devices)
    ms = []
    for _ in range(reps):
        t0 = time.perf_counter()
        model.predict_u8(imgs, k=5, devices=devices)
        ms.append((time.perf_counter() - t0) * 1000.0)
    ms.sort()
    return ms


def sweep():
    for g in ([n_gpus] if os.environ.get("SWEEP_ONLY_ALL_GPUS") == "1" else sorted({1, n_gpus})):
        ctx = b2.Context(tuple(range(g)))
        per_gpu = 4096 // g if g > 1 else 4096
        for name in ("resnet18", "alexnet"):
            ot = synth.ensure_ot(name, os.path.join(ROOT, "tests", "_cache", name + ".ot"))
            model = b2.Model(ctx, name, ot, max_batch=max(per_gpu, 1), micro_batch=256)
            pin, imgs = pinned_images(4096, seed=1)
            for bs in (1, 2, 4, 8, 16, 32, 64, 128, 256, 512, 1024, 2048, 4096):
                reps = 50 if bs <= 64 else (20 if bs <= 1024 else 8)
                ms = timed(model, imgs[:bs], reps)
                emit({"kind": "sweep", "model": name, "batch": bs, "gpus": g, "reps": reps,
                      "p50_ms": ms[len(ms) // 2], "p99_ms": ms[min(len(ms) - 1, int(0.99 * len(ms)))],
                      "images_per_s": bs / (sum(ms) / len(ms) / 1000.0), "api": "b2i_predict_u8 (blocking, pinned host buffers)"})
            if g == 1:
                rng = np.random.default_rng(3)
                raw = [rng.integers(0, 256, (375, 500, 3), dtype=np.uint8) for _ in range(256)]
                for bs in (1, 8, 32, 256):
                    for _ in range(2):
                        model.predict_raw(raw[:bs], k=5)
                    ms = []
                    for _ in range(10):
                        t0 = time.perf_counter()
                        model.predict_raw(raw[:bs], k=5)
                        ms.append((time.perf_counter() - t0) * 1000.0)
                    ms.sort()
                    emit({"kind": "sweep_raw", "model": name, "batch": bs, "gpus": 1, "image": "375x500 decoded u8", "p50_ms": ms[len(ms) // 2],
                          "images_per_s": bs / (sum(ms) / len(ms) / 1000.0), "api": "b2i_predict_raw_u8 (resize + crop + normalise on the GPU)"})
            model.close()
            pin.close()
        ctx.close()


def dual():
    if n_gpus < 2:
        emit({"kind": "dual", "skipped": "needs >= 2 GPUs, have %d" % n_gpus})
        return
    half = n_gpus // 2
    ctx = b2.Context(tuple(range(n_gpus)))
    mr = b2.Model(ctx, "resnet18", synth.ensure_ot("resnet18", os.path.join(ROOT, "tests", "_cache", "resnet18.ot")), max_batch=256)
    ma = b2.Model(ctx, "alexnet", synth.ensure_ot("alexnet", os.path.join(ROOT, "tests", "_cache", "alexnet.ot")), max_batch=256)
    pin, imgs = pinned_images(256 * half, seed=2)
    solo_r = timed(mr, imgs, 20, devices=(0, half))
    solo_a = timed(ma, imgs, 20, devices=(half, n_gpus - half))
    res = {}

    def job(key, model, devs):
        res[key] = timed(model, imgs, 20, devices=devs)

    th = [threading.Thread(target=job, args=("r", mr, (0, half))), threading.Thread(target=job, args=("a", ma, (half, n_gpus - half)))]
    for t in th:
        t.start()
    for t in th:
        t.join()
    rate = lambda ms: len(imgs) / (sum(ms) / len(ms) / 1000.0)
    emit({"kind": "dual", "gpus": n_gpus, "batch_per_job": len(imgs),
          "resnet18": {"gpus": [0, half], "alone_images_per_s": rate(solo_r), "together_images_per_s": rate(res["r"])},
          "alexnet": {"gpus": [half, n_gpus], "alone_images_per_s": rate(solo_a), "together_images_per_s": rate(res["a"])}})
    mr.close(); ma.close(); pin.close(); ctx.close()


if __name__ == "__main__":
    which = sys.argv[2] if len(sys.argv) > 2 else "both"
    if which in ("sweep", "both"):
        sweep()
    if which in ("dual", "both"):
        dual()
