"""Picks the committed subset of the reference's image fixtures and copies it under tests/golden/
(run from the repo root, in the container that holds /root/reference; then run make_golden.py):

    python tests/golden/select_fixtures.py [--dry-run]

The reference's only data for this path is test_files/imagenet_1k/train/<wnid>/<one JPEG> (1000
classes, reference src/services.rs:476,484-495).  All 1000 do not fit a source repository (131 MB), so
this script takes every image of each class of input SURVEY.md section 4 found in that set and a sample
of the ordinary ones, about 120 files / 15 MB:

  * every grayscale JPEG (the decoder must replicate the channel)
  * every image with a side shorter than 224 (the up-sampling filter)
  * every 4:2:2 image and the smallest 4:2:0 ones (chroma up-sampling in the decoder)
  * the widest, the tallest and the most extreme aspect ratios (long resampling plans, big crops)
  * square images (the `w * out_h == h * out_w` shortcut of the size rule)
  * the ordinary rest (4:4:4, landscape / portrait), smallest files first, up to the size budget

Selection is a pure function of the files (sorted by wnid, ties by name), so a second run picks the
same set.  The full 1000 are used by tools/parity_1000.py on the GPU box (pushed there once).
"""
import argparse
import os
import shutil
import sys

from PIL import Image, JpegImagePlugin

SRC = "/root/reference/test_files/imagenet_1k/train"
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
DST = os.path.join(ROOT, "tests", "golden", "imagenet_1k", "train")
BUDGET = 15 * 2 ** 20
TARGET = 120


def scan():
    rows = []
    for wnid in sorted(os.listdir(SRC)):
        files = sorted(os.listdir(os.path.join(SRC, wnid)))
        if not files:
            continue
        path = os.path.join(SRC, wnid, files[0])
        with Image.open(path) as im:
            w, h = im.size
            sampling = JpegImagePlugin.get_sampling(im) if im.format == "JPEG" else -1
            rows.append({"wnid": wnid, "file": files[0], "path": path, "w": w, "h": h, "gray": im.mode == "L",
                         "sampling": sampling, "bytes": os.path.getsize(path)})
    return rows


def select(rows):
    chosen, why = {}, {}

    def take(r, reason):
        if r["wnid"] not in chosen:
            chosen[r["wnid"]] = r
            why[r["wnid"]] = reason

    already = set(os.listdir(DST)) if os.path.isdir(DST) else set()
    for r in rows:
        if r["wnid"] in already:
            take(r, "committed earlier")
    for r in rows:
        if r["gray"]:
            take(r, "grayscale")
    for r in rows:
        if min(r["w"], r["h"]) < 224:
            take(r, "side < 224")
    for r in rows:
        if r["sampling"] == 1:
            take(r, "4:2:2")
    for r in sorted((r for r in rows if r["sampling"] == 2), key=lambda r: (r["bytes"], r["wnid"]))[:14]:
        take(r, "4:2:0")
    take(max(rows, key=lambda r: (r["w"], r["wnid"])), "widest")
    take(max(rows, key=lambda r: (r["h"], r["wnid"])), "tallest")
    for r in sorted(rows, key=lambda r: (-r["w"] / r["h"], r["wnid"]))[:3]:
        take(r, "aspect w/h %.2f" % (r["w"] / r["h"]))
    for r in sorted(rows, key=lambda r: (-r["h"] / r["w"], r["wnid"]))[:3]:
        take(r, "aspect h/w %.2f" % (r["h"] / r["w"]))
    for r in sorted((r for r in rows if r["w"] == r["h"]), key=lambda r: (r["bytes"], r["wnid"]))[:6]:
        take(r, "square")
    total = sum(r["bytes"] for r in chosen.values())
    for r in sorted(rows, key=lambda r: (r["bytes"], r["wnid"])):
        if len(chosen) >= TARGET or total + r["bytes"] > BUDGET:
            break
        if r["wnid"] not in chosen and r["sampling"] == 0 and not r["gray"] and r["w"] >= 224 and r["h"] >= 224:
            take(r, "ordinary")
            total += r["bytes"]
    return chosen, why


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--dry-run", action="store_true")
    args = ap.parse_args()
    if not os.path.isdir(SRC):
        sys.exit("needs the reference checkout at " + SRC)
    rows = scan()
    chosen, why = select(rows)
    total = sum(r["bytes"] for r in chosen.values())
    counts = {}
    for v in why.values():
        key = v.split(" ")[0] if v.startswith("aspect") else v
        counts[key] = counts.get(key, 0) + 1
    print("%d of %d images, %.1f MB: %s" % (len(chosen), len(rows), total / 2 ** 20, counts))
    if args.dry_run:
        return
    for wnid, r in sorted(chosen.items()):
        d = os.path.join(DST, wnid)
        os.makedirs(d, exist_ok=True)
        if not os.path.exists(os.path.join(d, r["file"])):
            shutil.copyfile(r["path"], os.path.join(d, r["file"]))
    with open(os.path.join(ROOT, "tests", "golden", "fixtures_manifest.tsv"), "w") as f:
        f.write("wnid\tfile\twidth\theight\tgray\tsampling\tbytes\twhy\n")
        for wnid, r in sorted(chosen.items()):
            f.write("%s\t%s\t%d\t%d\t%d\t%d\t%d\t%s\n" % (wnid, r["file"], r["w"], r["h"], r["gray"], r["sampling"],
                                                       r["bytes"], why[wnid]))


if __name__ == "__main__":
    main()
