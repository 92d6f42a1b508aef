"""numpy float32 restatement of tch's resize + centre crop (first half of
imagenet::load_image_and_resize(path, 224, 224), reference src/services.rs:492).

Target-size rule and crop: SURVEY.md Appendix B.1 / 8c R4 [tch-recall].  Resampling:
stb_image_resize's `stbir_resize_uint8` defaults (8c R3) [tch-recall] -- per axis Catmull-Rom when
the axis grows, Mitchell-Netravali (B = C = 1/3) evaluated in output space and re-normalised when it
shrinks (also at scale 1), edge clamp, linear u8 -> float (/255) -> filter -> *255 + 0.5 truncate.
All arithmetic is float32 with separate multiply and add; taps are accumulated in ascending input
order, horizontal pass first.  stb's source is not on this machine: parity with the real stb is
unpinned; the GPU kernels are checked bit-exactly against THIS restatement.
"""
import math

import numpy as np

f32 = np.float32


def resize_target(h, w):
    if w * 224 == h * 224:
        return 224, 224, 0, 0  # rw, rh, crop_x, crop_y
    ratio = max(224.0 / w, 224.0 / h)
    rw = max(int(ratio * w), 224)
    rh = max(int(ratio * h), 224)
    return rw, rh, (rw - 224) // 2, (rh - 224) // 2


def _catmullrom(x):
    x = np.abs(x).astype(f32)
    a = f32(1.0) - x * x * (f32(2.5) - f32(1.5) * x)
    b = f32(2.0) - x * (f32(4.0) + x * (f32(0.5) * x - f32(2.5)))
    return np.where(x < 1, a, np.where(x < 2, b, f32(0))).astype(f32)


def _mitchell(x):
    x = np.abs(x).astype(f32)
    a = (f32(16.0) + x * x * (f32(21.0) * x - f32(36.0))) / f32(18.0)
    b = (f32(32.0) + x * (f32(-60.0) + x * (f32(36.0) - f32(7.0) * x))) / f32(18.0)
    return np.where(x < 1, a, np.where(x < 2, b, f32(0))).astype(f32)


def plan_axis(in_size, out_size, crop):
    """For each of the 224 surviving outputs: (clamped input indices, float32 weights)."""
    scale = f32(out_size) / f32(in_size)
    taps = []
    if scale > 1:
        radius = f32(2.0) * scale
        for jj in range(224):
            centre = f32(jj + crop) + f32(0.5)
            lo = (centre - radius) / scale
            hi = (centre + radius) / scale
            in_centre = centre / scale
            first = int(math.floor(f32(lo + f32(0.5))))
            last = int(math.floor(f32(hi - f32(0.5))))
            pos = np.arange(first, last + 1)
            c = _catmullrom(in_centre - (pos.astype(f32) + f32(0.5)))
            while len(c) and c[0] == 0:      # stb drops leading zero taps
                c, pos = c[1:], pos[1:]
            total = f32(0)
            for v in c:
                total = f32(total + v)
            fs = f32(1.0) / total
            c = (c * fs).astype(f32)
            while len(c) and c[-1] == 0:
                c, pos = c[:-1], pos[:-1]
            taps.append((np.clip(pos, 0, in_size - 1), c))
        return taps
    radius = f32(2.0) / scale
    margin = int(math.ceil(f32(f32(4.0) / scale))) // 2
    n = np.arange(in_size + 2 * margin)
    centre = (n - margin).astype(f32) + f32(0.5)
    lo = (centre - radius) * scale
    hi = (centre + radius) * scale
    out_centre = centre * scale
    first = np.floor(lo + f32(0.5)).astype(np.int64)
    last = np.floor(hi - f32(0.5)).astype(np.int64)
    for jj in range(224):
        j = jj + crop
        sel = np.nonzero((first <= j) & (j <= last))[0]
        c = (_mitchell((f32(j) + f32(0.5)) - out_centre[sel]) * scale).astype(f32)
        # a contributor whose trailing coefficients are zero has its range cut short by stb, which
        # only removes zeros from the sums below -- adding them changes nothing
        total = f32(0)
        for v in c:
            total = f32(total + v)
        fs = f32(1.0) / total
        c = (c * fs).astype(f32)
        keep = c != 0
        taps.append((np.clip(sel[keep] - margin, 0, in_size - 1), c[keep]))
    return taps


def _apply(src, taps, axis):
    """src float32; resample `axis` to the 224 planned outputs, accumulating taps in order."""
    src = np.moveaxis(src, axis, 0)
    out = np.zeros((224,) + src.shape[1:], dtype=f32)
    for j, (idx, c) in enumerate(taps):
        acc = np.zeros(src.shape[1:], dtype=f32)
        for i, cv in zip(idx, c):
            acc = (acc + (src[i] * cv).astype(f32)).astype(f32)
        out[j] = acc
    return np.moveaxis(out, 0, axis)


def resize_crop_224(image_u8):
    """u8 [H,W,3] -> u8 [224,224,3]."""
    img = np.asarray(image_u8)
    h, w = img.shape[:2]
    rw, rh, cx, cy = resize_target(h, w)
    x = (img.astype(f32) / f32(255.0)).astype(f32)
    x = _apply(x, plan_axis(w, rw, cx), axis=1)   # horizontal first
    x = _apply(x, plan_axis(h, rh, cy), axis=0)
    x = np.clip(x, f32(0), f32(1)).astype(f32)
    return ((x * f32(255.0)).astype(f32) + f32(0.5)).astype(f32).astype(np.int32).astype(np.uint8)


def decode_image(path):
    """JPEG -> u8 [H,W,3] RGB (gray replicated), on the host.  The reference decodes with
    stb_image inside tch; PIL/libjpeg-turbo may differ by +-1-2 LSB, so parity is
    defined from the decoded pixels onward and both sides receive the same array."""
    from PIL import Image
    with Image.open(path) as im:
        return np.asarray(im.convert("RGB"))


def load_image_and_resize(path):
    return resize_crop_224(decode_image(path))
