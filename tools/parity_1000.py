#!/usr/bin/env python
"""The parity configurations BASELINE.json names, over ALL of the reference's fixture images:

    configs[0]  AlexNet   top-5 over test_files/imagenet_1k (1000 images), batch 32
    configs[1]  ResNet-18 over the same images, batch 256, bf16 and fp32 modes (plus fp16)

    python tools/parity_1000.py --images <dir with <wnid>/<file>.JPEG> --out profiles/rNN_parity_1000.json

The 1000 JPEGs live in the reference checkout (test_files/imagenet_1k/train, reference src/services.rs:476)
and are too big to commit; they are pushed to the GPU box once for this record (a git-ignored directory).
Per model / weight set / precision the record holds: the max-abs logit gap against the fp32 CPU restatement
in ABSOLUTE units and as a fraction of the logit standard deviation, how many images pass the margin filters
(top-1 / top-2 margin and 5th / 6th margin above the mode's tolerance) and how many top-1 / top-5-set
mismatches there are among them; the bit-exactness count of the GPU resize + crop against the restatement;
and the three ways from file to answer timed end to end (serial host decode, host thread pool, GPU decode).
Weights: the real pretrained files when they authenticate against the LFS pointers (oracle.synth.real_weights),
else the seeded synthetic sets -- the record says which.
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def load_crop(path):
    from oracle import stb_resize
    img = stb_resize.decode_image(path)
    return img.shape[0], img.shape[1], stb_resize.resize_crop_224(img)


def margins_report(ref, idx, tol, k=5):
    order = np.argsort(-ref, axis=1, kind="stable")
    srt = np.take_along_axis(ref, order, 1)
    out = {}
    for name, factor in (("margin_gt_tol", 1.0), ("margin_gt_2tol", 2.0)):
        m1 = (srt[:, 0] - srt[:, 1]) > factor * tol
        m5 = (srt[:, k - 1] - srt[:, k]) > factor * tol
        top1_bad = int(np.sum(m1 & (idx[:, 0] != order[:, 0])))
        set_bad = int(sum(1 for r in np.nonzero(m5)[0] if set(idx[r].tolist()) != set(order[r, :k].tolist())))
        out[name] = {"top1_compared": int(m1.sum()), "top1_mismatches": top1_bad,
                     "top5_set_compared": int(m5.sum()), "top5_set_mismatches": set_bad}
    out["top1_agreement_unfiltered"] = float(np.mean(idx[:, 0] == order[:, 0]))
    out["top5_set_agreement_unfiltered"] = float(np.mean([set(idx[r].tolist()) == set(order[r, :k].tolist())
                                                         for r in range(len(idx))]))
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--images", default=os.path.join(ROOT, "gpurun_in", "imagenet_1k", "train"))
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "parity_1000.json"))
    ap.add_argument("--limit", type=int, default=0)
    args = ap.parse_args()

    import torch
    import dmlc_b200 as b2
    from concurrent.futures import ProcessPoolExecutor
    from oracle import ot_format, synth, tch_oracle

    torch.set_num_threads(os.cpu_count() or 1)
    wnids = sorted(os.listdir(args.images))
    if args.limit:
        wnids = wnids[:args.limit]
    paths = [os.path.join(args.images, w, sorted(os.listdir(os.path.join(args.images, w)))[0]) for w in wnids]
    n = len(paths)
    t0 = time.perf_counter()
    with ProcessPoolExecutor(max_workers=min(32, os.cpu_count() or 1)) as ex:
        rows = list(ex.map(load_crop, paths, chunksize=8))
    crops = np.stack([r[2] for r in rows])
    rec = {"images": n, "source": args.images, "host_cores": os.cpu_count(),
           "oracle_preprocess_s": time.perf_counter() - t0,
           "sizes": {"min_side": int(min(min(r[0], r[1]) for r in rows)), "max_w": int(max(r[1] for r in rows)),
                     "max_h": int(max(r[0] for r in rows))}}
    labels = b2.load_labels(os.path.join(ROOT, "tests", "golden", "synset_words.txt"))
    class_of = {w: i for i, (w, _) in enumerate(labels)}
    truth = np.array([class_of.get(w, -1) for w in wnids])

    ctx = b2.Context((0,))
    cache = os.path.join(ROOT, "tests", "_cache")

    # ---------------- resize + crop: GPU kernels vs the restatement, every image
    from oracle import stb_resize
    exact = 0
    t0 = time.perf_counter()
    for p, c in zip(paths, crops):
        exact += int(np.array_equal(ctx.resize_crop(stb_resize.decode_image(p)), c))
    rec["resize_crop"] = {"bit_exact": exact, "of": n, "seconds": time.perf_counter() - t0}
    print("resize+crop bit-exact on %d / %d images" % (exact, n), flush=True)

    # ---------------- networks
    sets = []
    for name, arch, geoms in (("resnet18", "resnet18", [None]), ("resnet18_ps", "resnet18", [None]),
                              ("alexnet", "alexnet", ["tch", "torchvision"])):
        real = synth.real_weights(arch) if name != "resnet18_ps" else None
        if real:
            w = ot_format.read_ot(real)
            w = {k.replace("|", "."): v for k, v in w.items()}
            ot, kind = real, "pretrained (authenticated against the LFS pointer)"
        else:
            w = synth.weights(name)
            ot, kind = synth.ensure_ot(name, os.path.join(cache, name + ".ot")), "seeded synthetic"
        for g in geoms:
            sets.append((name, arch, g, w, ot, kind))
    rec["networks"] = []
    for name, arch, geom, w, ot, kind in sets:
        batch = 256 if arch == "resnet18" else 32
        t0 = time.perf_counter()
        kw = {"alexnet_geometry": geom} if geom else {}
        _, _, ref = tch_oracle.predict_u8(arch, w, crops, k=5, batch=64, **kw)
        cpu_s = time.perf_counter() - t0
        std, amax = float(ref.std()), float(np.abs(ref).max())
        entry = {"weights": name, "kind": kind, "geometry": geom, "batch": batch, "oracle_cpu_images_per_s": n / cpu_s,
                 "logit_std": std, "logit_absmax": amax, "precisions": {}}
        if (truth >= 0).all():
            entry["oracle_top1_accuracy_vs_directory_label"] = float(np.mean(np.argmax(ref, 1) == truth))
        for pname, prec, flat_tol in (("bf16", b2.PREC_BF16, 2e-2), ("fp16", b2.PREC_FP16, 2e-2), ("fp32", b2.PREC_FP32, 1e-4)):
            g = {"tch": b2.ALEXNET_TCH, "torchvision": b2.ALEXNET_TORCHVISION}.get(geom, b2.ALEXNET_TCH)
            with b2.Model(ctx, arch, ot, precision=prec, max_batch=batch, alexnet_geometry=g) as m:
                outs = [m.predict_u8(crops[lo:lo + batch], k=5, want_logits=True) for lo in range(0, n, batch)]
            idx, prob, logits = (np.concatenate([o[j] for o in outs]) for j in range(3))
            gap = float(np.abs(logits - ref).max())
            e = {"max_abs_gap": gap, "gap_over_std": gap / std, "flat_tolerance": flat_tol,
                 "meets_flat_tolerance": bool(gap <= flat_tol),
                 "meets_tolerance_in_units_of_std": bool(gap <= flat_tol * max(1.0, std)),
                 "filters_at_flat_tolerance": margins_report(ref, idx, flat_tol),
                 "filters_at_measured_gap": margins_report(ref, idx, gap)}
            if (truth >= 0).all():
                e["top1_accuracy_vs_directory_label"] = float(np.mean(idx[:, 0] == truth))
            entry["precisions"][pname] = e
            print("%-12s %-11s %s: gap %.3e abs = %.3f %% of std %.2f; flat %.0e %s; top-1 mismatches among margin>tol: %d of %d; "
                  "top-5 set: %d of %d" % (name, geom or "", pname, gap, 100 * gap / std, std, flat_tol,
                                           "MET" if gap <= flat_tol else "not met",
                                           e["filters_at_flat_tolerance"]["margin_gt_tol"]["top1_mismatches"],
                                           e["filters_at_flat_tolerance"]["margin_gt_tol"]["top1_compared"],
                                           e["filters_at_flat_tolerance"]["margin_gt_tol"]["top5_set_mismatches"],
                                           e["filters_at_flat_tolerance"]["margin_gt_tol"]["top5_set_compared"]), flush=True)
        rec["networks"].append(entry)

    # ---------------- from file names to answers: three decode arrangements, ResNet-18 bf16, batch 256
    ot = synth.real_weights("resnet18") or synth.ensure_ot("resnet18", os.path.join(cache, "resnet18.ot"))
    import shutil
    import tempfile
    wd = tempfile.mkdtemp()
    os.symlink(ot, os.path.join(wd, "resnet18.ot"))
    e2e = {}
    try:
        for mode in ("serial_pil", "pil_pool", "nvjpeg"):
            try:
                member = b2.MemberState(ctx, labels, weight_dir=wd, image_root=args.images, max_batch=256, models=("resnet18",),
                                        decoder="nvjpeg" if mode == "nvjpeg" else "pil",
                                        decode_threads=1 if mode == "serial_pil" else None)
            except b2.B2IError as ex:
                e2e[mode] = {"unavailable": str(ex)}
                continue
            try:
                sub = wnids if mode != "serial_pil" else wnids[:256]
                lab = [(w, dict(labels).get(w, "?")) for w in sub]
                leader = b2.Leader(member, lab)
                leader.resnet18.assigned_member_ids = [0]
                leader.run_job(leader.resnet18, batch=256)      # warm-up over the whole set: both staging lanes, decoder set-up, scratch growth
                leader.resnet18 = b2.Job("resnet18")
                leader.resnet18.assigned_member_ids = [0]
                t0 = time.perf_counter()
                leader.run_job(leader.resnet18, batch=256)
                dt = time.perf_counter() - t0
                rep = b2.jobs_report(leader.resnet18)
                durs = sorted(set(leader.resnet18.query_durations))
                e2e[mode] = {"images": rep["finished"], "seconds": dt, "images_per_s": rep["finished"] / dt,
                             "batch_ms_p50": 1000 * durs[len(durs) // 2] if durs else None,
                             "batch_ms_max": 1000 * durs[-1] if durs else None,
                             "threads": member.pool._max_workers, "decoder": member.decoder,
                             "accuracy_vs_directory_label": rep["accuracy"]}
                if mode == "nvjpeg":
                    e2e[mode]["backend"] = ctx.jpeg_backend()
                print("from file names, %-10s: %.0f images/s (%d images, %.2f s)" % (mode, rep["finished"] / dt, rep["finished"], dt), flush=True)
            finally:
                member.close()
        # decoder agreement: GPU-decoded vs host-decoded pixels through the same network
        if "images_per_s" in e2e.get("nvjpeg", {}):
            with b2.Model(ctx, "resnet18", ot, precision=b2.PREC_BF16, max_batch=256) as m:
                blobs = [open(p, "rb").read() for p in paths]
                ji, jst = [], []
                for lo in range(0, n, 256):
                    i, _, st = m.predict_jpeg(blobs[lo:lo + 256], k=5)
                    ji.append(i); jst.append(st)
                ji, jst = np.concatenate(ji), np.concatenate(jst)
                pi = np.concatenate([m.predict_u8(crops[lo:lo + 256], k=5)[0] for lo in range(0, n, 256)])
            e2e["nvjpeg"]["streams_refused"] = int((jst != 0).sum())
            e2e["nvjpeg"]["top1_agreement_with_host_decoded_pixels"] = float(np.mean(ji[jst == 0, 0] == pi[jst == 0, 0]))
    finally:
        shutil.rmtree(wd, ignore_errors=True)
    rec["from_file_names_resnet18_bf16_bs256"] = e2e
    if "images_per_s" in e2e.get("serial_pil", {}):
        for mode in ("pil_pool", "nvjpeg"):
            if "images_per_s" in e2e.get(mode, {}):
                e2e[mode]["speedup_over_serial_pil"] = e2e[mode]["images_per_s"] / e2e["serial_pil"]["images_per_s"]
    os.makedirs(os.path.dirname(os.path.abspath(args.out)), exist_ok=True)
    with open(args.out, "w") as f:
        json.dump(rec, f, indent=1)
    print("wrote", args.out)


if __name__ == "__main__":
    main()
