#!/usr/bin/env python
"""Throughput of the file-contents -> answers path (b2i_predict_jpeg) per decoder arrangement.

    python tools/jpeg_bench.py [--images DIR] [--batch 256] [--reps 3]

Runs itself once per arrangement in a child process (the decoder reads its environment switches once):
all backends, GPU Huffman only, hybrid with 8 / 16 / 32 host threads, pooled allocators off.  Prints one
JSON line per arrangement; B2I_TRACE=1 in the child prints the stage times of its last batch."""
import argparse
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def child(args):
    import numpy as np
    import dmlc_b200 as b2
    from oracle import synth
    wn = sorted(os.listdir(args.images))
    blobs = [open(os.path.join(args.images, w, sorted(os.listdir(os.path.join(args.images, w)))[0]), "rb").read() for w in wn]
    ot = synth.ensure_ot("resnet18", os.path.join(ROOT, "tests", "_cache", "resnet18.ot"))
    ctx = b2.Context((0,))
    with b2.Model(ctx, "resnet18", ot, max_batch=args.batch) as m:
        for lo in range(0, len(blobs), args.batch):      # warm-up: decoder set-up, buffers grow to their final size
            m.predict_jpeg(blobs[lo:lo + args.batch], k=5)
        if args.save:      # logits of every image, to compare the arrangements' decoded pixels through the network
            lg = np.concatenate([m.predict_jpeg(blobs[lo:lo + args.batch], k=5, want_logits=True)[2] for lo in range(0, len(blobs), args.batch)])
            np.save(args.save, lg)
        if args.callers > 1:   # several caller threads on one model: one request's decode overlaps another's forward
            import threading
            best = 0.0
            for r in range(args.reps):
                chunks = [blobs[lo:lo + args.batch] for lo in range(0, len(blobs), args.batch)] * 2
                def work(ix):
                    for c in chunks[ix::args.callers]:
                        m.predict_jpeg(c, k=5)
                th = [threading.Thread(target=work, args=(i,)) for i in range(args.callers)]
                t0 = time.perf_counter()
                [t.start() for t in th]
                [t.join() for t in th]
                best = max(best, 2 * len(blobs) / (time.perf_counter() - t0))
            print(json.dumps({"arrangement": args.label, "backend": ctx.jpeg_backend(), "callers": args.callers, "images_per_s": best,
                              "batch": args.batch, "images": 2 * len(blobs)}), flush=True)
            return
        best = 0.0
        for r in range(args.reps):
            t0 = time.perf_counter()
            bad = 0
            for lo in range(0, len(blobs), args.batch):
                if r == args.reps - 1 and lo + args.batch >= len(blobs):
                    os.environ["B2I_TRACE_NOW"] = "1"
                _, _, st = m.predict_jpeg(blobs[lo:lo + args.batch], k=5)
                bad += int((st != 0).sum())
            best = max(best, len(blobs) / (time.perf_counter() - t0))
        print(json.dumps({"arrangement": args.label, "backend": ctx.jpeg_backend(), "images_per_s": best, "refused": bad,
                          "batch": args.batch, "images": len(blobs), "mean_jpeg_bytes": sum(map(len, blobs)) // len(blobs)}), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--images", default=os.path.join(ROOT, "gpurun_in", "imagenet_1k", "train"))
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--label", default="")
    ap.add_argument("--callers", type=int, default=1)
    ap.add_argument("--save", default="")
    args = ap.parse_args()
    if args.label:
        return child(args)
    for label, env in (("all backends that open", {}),
                       ("GPU Huffman only", {"B2I_JPEG_BACKEND": "gpu"}),
                       ("hardware engines only", {"B2I_JPEG_BACKEND": "hardware"}),
                       ("hybrid, 8 host threads", {"B2I_JPEG_BACKEND": "cpu", "B2I_JPEG_THREADS": "8"}),
                       ("hybrid, 16 host threads", {"B2I_JPEG_BACKEND": "cpu", "B2I_JPEG_THREADS": "16"}),
                       ("hybrid, 32 host threads", {"B2I_JPEG_BACKEND": "cpu", "B2I_JPEG_THREADS": "32"}),
                       ("hybrid, 16 host threads, library allocators", {"B2I_JPEG_BACKEND": "cpu", "B2I_JPEG_THREADS": "16", "B2I_JPEG_NO_POOL": "1"}),
                       ("GPU Huffman only, library allocators", {"B2I_JPEG_BACKEND": "gpu", "B2I_JPEG_NO_POOL": "1"})):
        e = dict(os.environ, B2I_TRACE="1", **env)
        p = subprocess.run([sys.executable, os.path.abspath(__file__), "--images", args.images, "--batch", str(args.batch),
                            "--reps", str(args.reps), "--label", label], env=e, capture_output=True, text=True, timeout=600)
        print(p.stdout.strip() or json.dumps({"arrangement": label, "failed": p.stderr.strip()[-400:]}), flush=True)
        tr = [ln for ln in p.stderr.splitlines() if ln.startswith("[b2i trace]")]
        for ln in tr[-2:]:
            print("   ", ln, flush=True)


if __name__ == "__main__":
    main()
