"""Pins the oracle (CPU restatement of the tch path).  The reference holds no golden vectors for
this path (parity unpinned), so the anchors are: an independent implementation of
the same graphs (torchvision), the reference's layer/FLOP tables, and the committed golden vectors."""
import hashlib
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN, fixture_crops
from oracle import stb_resize, synth, tch_oracle

GV = np.load(os.path.join(GOLDEN, "golden_vectors.npz"))


def test_resnet18_graph_matches_torchvision():
    """Same weights through torchvision.models.resnet18 (an independent statement of the graph tch
    mirrors): logits agree to fp32 round-off."""
    import torchvision
    w = synth.resnet18_weights()
    m = torchvision.models.resnet18(num_classes=1000)
    sd = {k: v for k, v in w.items()}
    missing, unexpected = m.load_state_dict(sd, strict=False)
    assert not unexpected and all(k.endswith("num_batches_tracked") for k in missing)
    m.eval()
    x = tch_oracle.normalize_u8(synth.synthetic_images(2, seed=5))
    with torch.no_grad():
        ref, got = m(x), tch_oracle.resnet18_forward(w, x)
    assert torch.allclose(ref, got, rtol=0, atol=1e-4 * float(ref.abs().max()))


def test_alexnet_torchvision_geometry_matches_torchvision():
    import torchvision
    w = synth.alexnet_weights()
    m = torchvision.models.alexnet(num_classes=1000)
    m.load_state_dict(dict(w))
    m.eval()
    x = tch_oracle.normalize_u8(synth.synthetic_images(2, seed=6))
    with torch.no_grad():
        ref, got = m(x), tch_oracle.alexnet_forward(w, x, geometry="torchvision")
    assert torch.allclose(ref, got, rtol=0, atol=1e-4 * float(ref.abs().max()))


def test_alexnet_tch_geometry_shapes():
    """features.3 = k5 p1 s2: 27 -> 13 -> 6 -> 6 -> 6 -> 2, then adaptive pool 2x2 -> 6x6 replicates
    (right column)."""
    w = synth.alexnet_weights()
    taps = []
    with torch.no_grad():
        tch_oracle.alexnet_forward(w, tch_oracle.normalize_u8(synth.synthetic_images(1)), "tch", taps)
    shapes = {n: tuple(t.shape[1:]) for n, t in taps}
    assert shapes["features.0"] == (64, 55, 55) and shapes["pool0"] == (64, 27, 27)
    assert shapes["features.3"] == (192, 13, 13) and shapes["pool1"] == (192, 6, 6)
    assert shapes["features.10"] == (256, 6, 6) and shapes["pool2"] == (256, 2, 2)
    assert shapes["avgpool"] == (256, 6, 6)
    ap = dict(taps)["avgpool"][0, 0]
    p2 = dict(taps)["pool2"][0, 0]
    assert torch.equal(ap, p2.repeat_interleave(3, 0).repeat_interleave(3, 1))


def test_resnet18_mac_count_matches_survey():
    """1,814,073,344 MAC per image over conv + FC."""
    w = synth.resnet18_weights()
    macs = 0
    hw = {"conv1": 112, "layer1": 56, "layer2": 28, "layer3": 14, "layer4": 7}
    for name, t in w.items():
        if name.endswith("weight") and t.dim() == 4:
            macs += t.numel() * hw[name.split(".")[0]] ** 2
    macs += w["fc.weight"].numel()
    assert macs == 1_814_073_344


def test_normalize_matches_definition():
    img = synth.synthetic_images(1, seed=2)
    x = tch_oracle.normalize_u8(img)[0].numpy()
    for c in range(3):
        want = (img[0, :, :, c].astype(np.float32) / np.float32(255.0) - np.float32(tch_oracle.MEAN[c])) / np.float32(tch_oracle.STD[c])
        assert np.allclose(x[c], want, rtol=0, atol=1e-6)


def test_stem_fma_normalise_is_exact_in_bf16():
    """csrc/stem.cu normalises with one fp32 FMA per byte (u * a + b, the constants below) instead of
    the oracle's divide / subtract / divide.  For every one of the 3 x 256 possible inputs both give
    the same bf16 value, so the fused stem sees bit-identical pixels."""
    na = [float.fromhex("0x1.18926cp-6"), float.fromhex("0x1.1ed5b2p-6"), float.fromhex("0x1.1d8f56p-6")]
    nb = [float.fromhex("-0x1.0f177ap+1"), float.fromhex("-0x1.04924ap+1"), float.fromhex("-0x1.cdf012p+0")]
    u = np.arange(256, dtype=np.float32)
    for c in range(3):
        assert np.float32(na[c]) == np.float32(1.0 / (255.0 * tch_oracle.STD[c]))
        assert np.float32(nb[c]) == np.float32(-tch_oracle.MEAN[c] / tch_oracle.STD[c])
        exact = ((u / np.float32(255.0)) - np.float32(tch_oracle.MEAN[c])) / np.float32(tch_oracle.STD[c])
        fma = (u.astype(np.float64) * na[c] + nb[c]).astype(np.float32)   # the product and sum are exact in f64
        to_bf16 = lambda a: torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32)).to(torch.bfloat16)
        assert torch.equal(to_bf16(exact), to_bf16(fma))


def test_softmax_topk_sorted_and_normalised():
    lg = torch.randn(7, 1000, generator=torch.Generator().manual_seed(0))
    idx, prob = tch_oracle.softmax_topk(lg, 5)
    assert idx.shape == (7, 5) and (prob[:, :-1] >= prob[:, 1:]).all()
    assert torch.equal(idx[:, 0].long(), lg.argmax(1))


def test_resize_target_rule():
    assert stb_resize.resize_target(332, 500) == (337, 224, 56, 0)      # the model_utils.rs:8 sample image
    assert stb_resize.resize_target(500, 331) == (224, 338, 0, 57)
    assert stb_resize.resize_target(100, 100) == (224, 224, 0, 0)       # square shortcut
    assert stb_resize.resize_target(94, 120) == (285, 224, 30, 0)       # up-sampling path
    rw, rh, cx, cy = stb_resize.resize_target(2877, 100)
    assert rw == 224 and rh >= 224 and cx == 0 and cy == (rh - 224) // 2


def test_resize_identity_like_properties():
    flat = np.full((300, 400, 3), 77, np.uint8)
    assert (stb_resize.resize_crop_224(flat) == 77).all()                # weights sum to one on every output
    ramp = np.tile(np.arange(224, dtype=np.uint8)[None, :, None], (448, 1, 3))
    out = stb_resize.resize_crop_224(ramp)                               # width already 224 (scale 1): Mitchell blur of a ramp is the ramp
    assert np.abs(out[:, 4:-4, :].astype(int) - ramp[:224, 4:-4, :].astype(int)).max() <= 1


def test_golden_vectors_still_reproduce():
    crops = fixture_crops()
    assert [hashlib.sha256(c.tobytes()).hexdigest() for c in crops] == list(GV["crop_sha256"])
    i = list(GV["wnids"]).index("n03825788")
    assert np.array_equal(crops[i], GV["crop_n03825788"])
    sel = list(range(0, len(crops), 9))                      # a spread of the set keeps the CPU suite short
    _, _, lg = tch_oracle.predict_u8("resnet18", synth.resnet18_weights(), crops[sel])
    assert np.allclose(lg, GV["resnet18_logits"][sel], rtol=0, atol=2e-4 * np.abs(lg).max())
    _, _, lp = tch_oracle.predict_u8("resnet18", synth.weights("resnet18_ps"), crops[sel])
    assert np.allclose(lp, GV["resnet18_ps_logits"][sel], rtol=0, atol=2e-4 * np.abs(lp).max())
    assert 2.0 < float(GV["resnet18_ps_logits"].std()) < 3.0    # "pretrained scale": what flat tolerances assume
    _, _, la = tch_oracle.predict_u8("alexnet", synth.alexnet_weights(), crops[:3], alexnet_geometry="tch")
    assert np.allclose(la, GV["alexnet_tch_logits"][:3], rtol=0, atol=2e-4 * np.abs(la).max())


def test_fixture_set_covers_the_input_classes_of_the_reference_set():
    """The committed subset holds every kind of file SURVEY.md section 4 found among the reference's 1000:
    grayscale, sides under 224 (up-sampling), 4:2:2 and 4:2:0 chroma, extreme aspect ratios, squares."""
    rows = [ln.rstrip("\n").split("\t") for ln in open(os.path.join(GOLDEN, "fixtures_manifest.tsv"))][1:]
    assert len(rows) == len(GV["wnids"]) >= 100
    w = np.array([int(r[2]) for r in rows]); h = np.array([int(r[3]) for r in rows])
    gray = np.array([int(r[4]) for r in rows]); samp = np.array([int(r[5]) for r in rows])
    assert gray.sum() >= 19 and (np.minimum(w, h) < 224).sum() >= 18
    assert (samp == 1).sum() >= 9 and (samp == 2).sum() >= 10
    assert w.max() == 2304 and h.max() == 2877 and (w == h).sum() >= 6
    assert (w / h).max() > 2.5 and (h / w).max() > 2.0


def test_real_weight_hook_authenticates_by_size_and_sha256(tmp_path):
    """oracle.synth.real_weights: only a file with the LFS pointer's size and sha256 is taken (reference
    pretrained_models/*.ot:2-3); pointers, truncated files and other weights are ignored."""
    assert synth.LFS["resnet18"] == ("f349134076ce5f2dc50f1003b798fe068c7ab0695c757450723d259441b20ebb", 46831783)
    assert synth.LFS["alexnet"] == ("8fb43520a45c164aba438753cdc5fd4b6f5c5dd5436db6d69efdf1e5532ff0e8", 244408272)
    (tmp_path / "resnet18.ot").write_text("version https://git-lfs.github.com/spec/v1\noid sha256:f349\nsize 46831783\n")
    assert synth.real_weights("resnet18", [str(tmp_path)]) is None
    blob = b"x" * 1000
    import hashlib
    saved = dict(synth.LFS)
    try:
        synth.LFS["resnet18"] = (hashlib.sha256(blob).hexdigest(), len(blob))
        (tmp_path / "resnet18.ot").write_bytes(blob)
        assert synth.real_weights("resnet18", [str(tmp_path)]) == str(tmp_path / "resnet18.ot")
        (tmp_path / "resnet18.ot").write_bytes(b"y" * 1000)          # right size, wrong content
        assert synth.real_weights("resnet18", [str(tmp_path)]) is None
    finally:
        synth.LFS.update(saved)


def test_unknown_model_is_an_error():
    with pytest.raises(ValueError):
        tch_oracle.forward("vgg16", {}, torch.zeros(1, 3, 224, 224))
