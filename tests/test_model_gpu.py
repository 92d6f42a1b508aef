"""Whole-network parity on the GPU through the C ABI (b2i_model_load / b2i_predict_u8) against the
oracle and the committed golden vectors: logits, softmax probabilities and top-5 indices, for both
models, three precisions, both AlexNet geometries, ragged / empty / single-image batches.

Fixture set: the 120 JPEGs tests/golden/select_fixtures.py picked from the reference's
test_files/imagenet_1k (every grayscale, every side < 224, every 4:2:2, extreme aspect ratios ...), run at
the batch sizes BASELINE.json names: ResNet-18 at batch 256 (the set cycled to fill it), AlexNet at batch 32.
The whole 1000 go through the same checks in tools/parity_1000.py (record under profiles/).

Tolerances (BASELINE.json north_star): fp32 mode 1e-4 max-abs, bf16 mode 2e-2 max-abs on logits.  Both are
ABSOLUTE figures meant for logits of pretrained scale (standard deviation 2-3).  Two weight sets:
  * "ps" (oracle.synth resnet18_ps: logits of std 2.5, |max| 13) and AlexNet (std 2.1): the bounds are applied
    FLAT.  fp32 mode and fp16 mode (PREC_FP16, the tensor-core kernels with IEEE-half storage) meet them;
    bf16 mode does not and is not claimed to: its distance is printed and bounded by 4 % of the logit std.
  * the default synthetic ResNet-18 (std 8.9, |max| 47): the bounds are applied in units of the logit std.
bf16 mode, two separate statements:
  (1) kernel exactness: against the oracle evaluated WITH bf16 storage (oracle.tch_oracle.*_bf16:
      same roundings, fp32 sums) the gap must be within 2.5e-2 units;
  (2) cost of the dtype: against the fp32 oracle the kernel may be no further away than 1.25x the
      distance of that bf16 evaluation itself (measured 2.4-3.5 % of the logit std, the same
      2-4 % SURVEY.md Appendix E found for torch's own bf16 CPU path).
Top-5: on every image whose top-1 / top-2 margin exceeds twice the tolerance the top-1 index is identical;
where the 5th / 6th margin does too, the top-5 SET is identical; where every neighbouring pair among the first
six is that far apart, the order is identical.  The counts of images passing each filter are printed."""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN, fixture_crops, fixture_paths

pytestmark = pytest.mark.gpu

BF16, FP32, FP16 = 0, 1, 2
GV = np.load(os.path.join(GOLDEN, "golden_vectors.npz"))
NFIX = len(GV["wnids"])


def tol_for(precision, ref_logits, flat=False):
    unit = 1.0 if flat else max(1.0, float(np.std(ref_logits)))
    return (1e-4 if precision == FP32 else 2e-2) * unit


def bf16_budget(ref_logits):
    """Distance a bf16 evaluation may have from the fp32 oracle: 4 % of the logit std (measured over the 120
    fixtures: 2.4 % ResNet-18, 3.3-3.5 % AlexNet; over all 1000 reference images 2.7 % / 3.5 %,
    profiles/r03_parity_1000.json; see the module docstring)."""
    return 4e-2 * max(1.0, float(np.std(ref_logits)))


def check_topk(ref_logits, idx, tol, what):
    """The margin-filtered top-k comparison of the module docstring; returns the filter counts."""
    order = np.argsort(-ref_logits, axis=1, kind="stable")
    srt = np.take_along_axis(ref_logits, order, 1)
    k = idx.shape[1]
    n1 = nset = nord = 0
    for r in range(ref_logits.shape[0]):
        margins = srt[r, :k] - srt[r, 1:k + 1]
        if margins[0] > 2 * tol:
            n1 += 1
            assert idx[r, 0] == order[r, 0], (what, "top-1", r, idx[r], order[r, :k])
        if margins[k - 1] > 2 * tol:
            nset += 1
            assert set(idx[r].tolist()) == set(order[r, :k].tolist()), (what, "top-%d set" % k, r, idx[r], order[r, :k])
        if (margins > 2 * tol).all():
            nord += 1
            assert np.array_equal(idx[r], order[r, :k]), (what, "order", r, idx[r], order[r, :k])
    print("%s: of %d images, top-1 compared on %d, top-%d set on %d, full order on %d (margin > %.3e): all identical" %
          (what, ref_logits.shape[0], n1, k, nset, nord, 2 * tol))
    return n1, nset, nord


def check_against(ref_logits, idx, prob, logits, precision, what, emu_logits=None, flat=False):
    tol = tol_for(precision, ref_logits, flat)
    gap = float(np.abs(logits - ref_logits).max())
    std = float(np.std(ref_logits))
    print("%s: logits max-abs gap vs fp32 oracle %.3e absolute = %.3f %% of the logit std %.3f (|max| %.2f); bound %.3e %s" %
          (what, gap, 100.0 * gap / std, std, np.abs(ref_logits).max(), tol, "FLAT" if flat else "in units of max(1, std)"))
    if precision == FP32 or precision == FP16:
        # fp16 storage (11-bit significands) meets the 2e-2 bar against the fp32 oracle OUTRIGHT, no dtype allowance
        assert gap <= tol, what
    else:
        tol = tol_for(precision, ref_logits, False)       # bf16 storage: never claimed flat
        gap_emu = float(np.abs(logits - emu_logits).max())
        emu_vs_fp32 = float(np.abs(emu_logits - ref_logits).max())
        print("%s: vs bf16-evaluated oracle %.3e (bound %.3e); that oracle vs fp32 oracle %.3e; flat 2e-2 vs fp32 oracle: %s" %
              (what, gap_emu, tol, emu_vs_fp32, "met" if gap <= 2e-2 else "NOT met (bf16 storage)"))
        # (1) kernel exactness.  The CPU evaluation rounds where the kernels round, but sums in another order and
        # does not fuse what the stems fuse: over the 120-image set the two differ by up to 2.05 % of the logit
        # std (AlexNet), so the bound is 2.5e-2 units here
        assert gap_emu <= 1.25 * tol, what
        assert gap <= 1.25 * emu_vs_fp32 + 1e-3, what    # (2) no worse than bf16 storage itself
        assert gap <= bf16_budget(ref_logits), what
        tol = max(tol, 1.25 * emu_vs_fp32)                # rank comparisons below use the dtype's real budget
    # softmax + top-5 of the kernel's own logits must be what torch computes from them
    t = torch.from_numpy(logits)
    p = torch.softmax(t, -1)
    pv, pi = torch.topk(p, idx.shape[1], -1)
    assert np.abs(prob - pv.numpy()).max() <= 1e-5
    # top-k indices vs the ORACLE's logits wherever ranks are separated by more than the tolerance
    n1, nset, nord = check_topk(ref_logits, idx, tol, what)
    assert n1 >= ref_logits.shape[0] // 2, (what, "the margin filter left too little to compare", n1)


def resnet18_batch256(crops):
    """BASELINE.json configs[1]: ResNet-18 at batch 256 -- the fixture set cycled to fill the batch."""
    rep = np.resize(np.arange(len(crops)), 256)
    return crops[rep], rep


@pytest.mark.parametrize("precision", [BF16, FP32, FP16], ids=["bf16", "fp32", "fp16"])
def test_resnet18_fixture_images_match_golden(pkg, ctx, resnet18_ot, precision):
    crops = fixture_crops()
    batch, rep = resnet18_batch256(crops)
    with pkg.Model(ctx, "resnet18", resnet18_ot, precision=precision, max_batch=256) as m:
        idx, prob, logits = m.predict_u8(batch, k=5, want_logits=True)
    # a repeated image gets the same answer wherever it sits in the batch
    assert np.array_equal(logits[NFIX:], logits[rep[NFIX:]])
    idx, prob, logits = idx[:NFIX], prob[:NFIX], logits[:NFIX]
    emu = None
    if precision == BF16:
        from oracle import synth, tch_oracle
        with torch.no_grad():
            emu = tch_oracle.resnet18_forward_bf16(synth.resnet18_weights(), crops).numpy()
    check_against(GV["resnet18_logits"], idx, prob, logits, precision, "resnet18 fixtures bs256", emu)


@pytest.mark.parametrize("precision", [BF16, FP32, FP16], ids=["bf16", "fp32", "fp16"])
def test_resnet18_pretrained_scale_logits_flat_bounds(pkg, ctx, resnet18_ps_ot, precision):
    """The north-star tolerances as written -- max-abs 2e-2 (16-bit modes) / 1e-4 (fp32 mode), no scaling --
    on the weight set whose logits have pretrained scale.  fp32 and fp16 modes must meet them; for bf16 the
    absolute gap is printed and held to 4 % of the logit std (it does not meet a flat 2e-2: 8 mantissa bits)."""
    crops = fixture_crops()
    batch, rep = resnet18_batch256(crops)
    with pkg.Model(ctx, "resnet18", resnet18_ps_ot, precision=precision, max_batch=256) as m:
        idx, prob, logits = m.predict_u8(batch, k=5, want_logits=True)
    idx, prob, logits = idx[:NFIX], prob[:NFIX], logits[:NFIX]
    emu = None
    if precision == BF16:
        from oracle import synth, tch_oracle
        with torch.no_grad():
            emu = tch_oracle.resnet18_forward_bf16(synth.weights("resnet18_ps"), crops).numpy()
    check_against(GV["resnet18_ps_logits"], idx, prob, logits, precision, "resnet18 (pretrained-scale logits) bs256", emu,
                  flat=True)


@pytest.mark.parametrize("geometry", [0, 1], ids=["tch", "torchvision"])
@pytest.mark.parametrize("precision", [BF16, FP32, FP16], ids=["bf16", "fp32", "fp16"])
def test_alexnet_fixture_images_match_golden(pkg, ctx, alexnet_ot, precision, geometry):
    """BASELINE.json configs[0]: AlexNet at batch 32 (three full batches and one of 24); AlexNet's synthetic
    logits have std 2.1, so the bounds are the flat ones."""
    crops = fixture_crops()
    outs = []
    with pkg.Model(ctx, "alexnet", alexnet_ot, precision=precision, max_batch=32, alexnet_geometry=geometry) as m:
        for lo in range(0, NFIX, 32):
            outs.append(m.predict_u8(crops[lo:lo + 32], k=5, want_logits=True))
    idx, prob, logits = (np.concatenate([o[j] for o in outs]) for j in range(3))
    ref = GV["alexnet_tch_logits" if geometry == 0 else "alexnet_torchvision_logits"]
    emu = None
    if precision == BF16:
        from oracle import synth, tch_oracle
        with torch.no_grad():
            emu = tch_oracle.alexnet_forward_bf16(synth.alexnet_weights(), crops, "tch" if geometry == 0 else "torchvision").numpy()
    check_against(ref, idx, prob, logits, precision, "alexnet fixtures bs32 geometry %d" % geometry, emu, flat=True)


@pytest.mark.parametrize("conv_mode", [0, 1], ids=["gather", "im2col"])
def test_resnet18_per_layer_parity_bf16(pkg, ctx, resnet18_ot, conv_mode):
    """Every activation of the bf16 path against an oracle that rounds at the same places
    (bf16 weights after BN fold are NOT emulated: per-layer bound is 3% of the layer's scale)."""
    from oracle import synth, tch_oracle
    img = synth.synthetic_images(3, seed=4)
    w = synth.resnet18_weights()
    taps = []
    with torch.no_grad():
        tch_oracle.resnet18_forward(w, tch_oracle.normalize_u8(img), taps)
    with pkg.Model(ctx, "resnet18", resnet18_ot, precision=BF16, max_batch=3, keep_tensors=True, conv_mode=conv_mode,
                   fused_stem=False) as m:
        m.predict_u8(img, k=5)
        shapes = [m.tensor_shape(t) for t in range(m.num_tensors())]
        # tensors in execution order: 0 input, 1 conv1, 2 maxpool, then per block (downsample?), conv1, conv2
        got = {t: m.read_tensor(t, 3) for t in range(m.num_tensors() - 1)}
    names = {"conv1": 1, "maxpool": 2}
    tid = 3
    for L in range(1, 5):
        for B in range(2):
            tid += (3 if (L > 1 and B == 0) else 2)
            names["layer%d.%d" % (L, B)] = tid - 1
    names["avgpool"] = tid
    worst = 0.0
    for name, t in taps:
        ref = t.permute(0, 2, 3, 1).numpy()
        g = got[names[name]]
        assert g.shape == ref.shape, (name, g.shape, ref.shape, shapes[names[name]])
        rel = float(np.abs(g - ref).max() / max(1e-6, np.abs(ref).max()))
        print("%-10s tensor %2d max-abs/scale %.4f" % (name, names[name], rel))
        worst = max(worst, rel)
        assert rel < 0.03, name
    assert worst > 0  # the comparison really ran


def test_fused_stem_matches_unfused_and_oracle(pkg, ctx, resnet18_ot):
    """normalise + conv1 + bn1 + relu + max-pool as ONE kernel: its pooled map against the oracle's
    max-pool output and against the three-kernel path (same roundings, different K order)."""
    from oracle import synth, tch_oracle
    img = synth.synthetic_images(5, seed=8)
    img[0, :7] = 0
    img[1, :, -9:] = 255   # saturated borders exercise the zero-padding edges
    taps = []
    with torch.no_grad():
        tch_oracle.resnet18_forward(synth.resnet18_weights(), tch_oracle.normalize_u8(img), taps)
    ref = dict(taps)["maxpool"].permute(0, 2, 3, 1).numpy()
    with pkg.Model(ctx, "resnet18", resnet18_ot, precision=BF16, max_batch=5, keep_tensors=True, fused_stem=True) as m:
        _, _, lg_f = m.predict_u8(img, k=5, want_logits=True)
        assert m.tensor_shape(0) == (56, 56, 64)
        fused = m.read_tensor(0, 5)
    with pkg.Model(ctx, "resnet18", resnet18_ot, precision=BF16, max_batch=5, keep_tensors=True, fused_stem=False) as m:
        _, _, lg_u = m.predict_u8(img, k=5, want_logits=True)
        unfused = m.read_tensor(2, 5)
    scale = float(np.abs(ref).max())
    print("stem: fused vs oracle %.4f, unfused vs oracle %.4f, fused vs unfused %.4f (of scale %.2f)" % (
        np.abs(fused - ref).max() / scale, np.abs(unfused - ref).max() / scale, np.abs(fused - unfused).max() / scale, scale))
    assert np.abs(fused - ref).max() <= 0.01 * scale
    assert np.abs(fused - unfused).max() <= 2.0 ** -7 * scale      # at most one bf16 ulp flip at the top of the range
    assert np.abs(lg_f - lg_u).max() <= tol_for(BF16, lg_u)


def test_alexnet_fused_stem_matches_unfused(pkg, ctx, alexnet_ot):
    """normalise + features.0 + ReLU as ONE kernel (two-plane 4-byte pixels, descriptors straight into the input
    patch): its 55 x 55 x 64 map against the two-kernel path (same roundings, different K order), and the logits
    of both paths against each other; zeroed / saturated borders exercise the zero-padding edges."""
    from oracle import synth
    img = synth.synthetic_images(6, seed=18)
    img[0, :5] = 0
    img[1, :, -7:] = 255
    img[2, -3:, :] = 255
    with pkg.Model(ctx, "alexnet", alexnet_ot, precision=BF16, max_batch=6, keep_tensors=True, fused_stem=True) as m:
        _, _, lg_f = m.predict_u8(img, k=5, want_logits=True)
        assert m.tensor_shape(0) == (55, 55, 64)
        fused = m.read_tensor(0, 6)
    with pkg.Model(ctx, "alexnet", alexnet_ot, precision=BF16, max_batch=6, keep_tensors=True, fused_stem=False) as m:
        _, _, lg_u = m.predict_u8(img, k=5, want_logits=True)
        assert m.tensor_shape(1) == (55, 55, 64)
        unfused = m.read_tensor(1, 6)
    scale = float(np.abs(unfused).max())
    print("alexnet stem: fused vs unfused %.5f of scale %.2f" % (np.abs(fused - unfused).max() / scale, scale))
    assert np.abs(fused - unfused).max() <= 2.0 ** -7 * scale      # at most one bf16 ulp flip at the top of the range
    assert np.abs(lg_f - lg_u).max() <= tol_for(BF16, lg_u)


def test_kernel_variants_agree(pkg, ctx, resnet18_ot):
    """Every optional kernel path (fused stem, halo patch, CTA pairs, fused downsample) against the
    plain persistent im2col configuration: same roundings, so logits agree to bf16 noise."""
    from oracle import synth
    img = synth.synthetic_images(6, seed=61)
    def run(**kw):
        with pkg.Model(ctx, "resnet18", resnet18_ot, precision=BF16, max_batch=6, **kw) as m:
            return m.predict_u8(img, k=5, want_logits=True)[2]
    plain = run(fused_stem=False, patch_conv=0, pair_mma=0, fuse_downsample=0)
    tol = 0.5 * tol_for(BF16, plain)
    for kw in (dict(fuse_downsample=1, patch_conv=0, pair_mma=0, fused_stem=False), dict(pair_mma=128, fuse_downsample=1),
               dict(pair_mma=64, fuse_downsample=0), dict(), dict(conv_mode=2, pair_mma=0, fuse_downsample=0)):
        got = run(**kw)
        gap = float(np.abs(got - plain).max())
        print("variant %s: gap to plain path %.3e (bound %.3e)" % (kw, gap, tol))
        assert gap <= tol, kw
    # the fused downsample writes the same tensor the separate kernel would
    with pkg.Model(ctx, "resnet18", resnet18_ot, precision=BF16, max_batch=6, keep_tensors=True, fused_stem=False, patch_conv=0,
                   pair_mma=0, fuse_downsample=0) as m0, \
            pkg.Model(ctx, "resnet18", resnet18_ot, precision=BF16, max_batch=6, keep_tensors=True, fused_stem=False, patch_conv=0,
                      pair_mma=0, fuse_downsample=1) as m1:
        m0.predict_u8(img, k=5)
        m1.predict_u8(img, k=5)
        for tid in (7, 12, 17):   # layer2/3/4 .0 downsample outputs (tensor order: see the per-layer test)
            assert np.array_equal(m0.read_tensor(tid, 6), m1.read_tensor(tid, 6)), tid


def test_batch_shapes_ragged_empty_single(pkg, ctx, resnet18_ot):
    from oracle import synth
    img = synth.synthetic_images(11, seed=0)
    with pkg.Model(ctx, "resnet18", resnet18_ot, precision=BF16, max_batch=16, micro_batch=4) as m:
        idx_all, prob_all, lg_all = m.predict_u8(img, k=5, want_logits=True)        # 4+4+3: ragged micro-batches
        idx0, prob0 = m.predict_u8(img[:0], k=5)
        assert idx0.shape == (0, 5) and prob0.shape == (0, 5)
        idx1, prob1, lg1 = m.predict_u8(img[4:5], k=5, want_logits=True)
        # an image's result does not depend on its batch position or on its neighbours
        assert np.array_equal(lg1[0], lg_all[4]) and np.array_equal(idx1[0], idx_all[4])
        with pytest.raises(pkg.B2IError):
            m.predict_u8(synth.synthetic_images(17, seed=1), k=5)                  # > max_batch
        with pytest.raises(pkg.B2IError):
            m.predict_u8(img, k=9)
    ref = GV["resnet18_logits_synthetic4"]
    assert np.abs(lg_all[:4] - ref).max() <= bf16_budget(ref)


def test_submit_wait_overlaps_and_matches_blocking_predict(pkg, ctx, resnet18_ot):
    """b2i_submit_u8 / b2i_wait: two requests in flight give the same answers as blocking calls, a
    third is refused, tickets are checked."""
    from oracle import synth
    a, b = synth.synthetic_images(6, seed=21), synth.synthetic_images(3, seed=22)
    with pkg.Model(ctx, "resnet18", resnet18_ot, precision=BF16, max_batch=8, micro_batch=4) as m:
        ia, pa = m.predict_u8(a, k=5)
        ib, pb = m.predict_u8(b, k=5)
        oa = (np.empty((6, 5), np.int32), np.empty((6, 5), np.float32))
        ob = (np.empty((3, 5), np.int32), np.empty((3, 5), np.float32))
        ta = m.submit_u8(a, *oa)
        tb = m.submit_u8(b, *ob)
        with pytest.raises(pkg.B2IError):
            m.submit_u8(a, np.empty((6, 5), np.int32), np.empty((6, 5), np.float32))   # third in flight
        with pytest.raises(pkg.B2IError):
            m.predict_u8(a, k=5)                                                        # blocking call while tickets are open
        m.wait(ta)
        with pytest.raises(pkg.B2IError):
            m.wait(ta)                                                                  # already collected
        m.wait(tb)
        assert np.array_equal(oa[0], ia) and np.array_equal(oa[1], pa)
        assert np.array_equal(ob[0], ib) and np.array_equal(ob[1], pb)
        for r in range(4):                                                              # steady-state pipeline
            t = m.submit_u8(a, *oa)
            m.wait(t)
        assert np.array_equal(oa[0], ia)


def test_graph_replay_of_small_batches_is_transparent(pkg, ctx, resnet18_ot, alexnet_ot):
    """Small forwards are captured into a CUDA graph on their second use and replayed afterwards:
    the first (eager), second (captured) and later (replayed) calls must return identical results,
    also when other batch sizes are interleaved."""
    from oracle import synth
    imgs = synth.synthetic_images(9, seed=41)
    for name, ot in (("resnet18", resnet18_ot), ("alexnet", alexnet_ot)):
        with pkg.Model(ctx, name, ot, precision=BF16, max_batch=16) as m:
            first = {n: m.predict_u8(imgs[:n], k=5, want_logits=True) for n in (1, 3, 9)}
            for rep in range(4):
                for n in (3, 1, 9):
                    i, p, lg = m.predict_u8(imgs[:n], k=5, want_logits=True)
                    assert np.array_equal(lg, first[n][2]) and np.array_equal(i, first[n][0]), (name, n, rep)
            # a different image through a captured shape must not see stale inputs
            other = synth.synthetic_images(3, seed=42)
            _, _, lo = m.predict_u8(other, k=5, want_logits=True)
            assert not np.array_equal(lo, first[3][2])
            _, _, lo2 = m.predict_u8(other, k=5, want_logits=True)
            assert np.array_equal(lo, lo2)


def test_predict_raw_equals_resize_then_predict(pkg, ctx, resnet18_ot):
    """b2i_predict_raw_u8 (decoded images of any size in, top-k out) is bit-identical to
    b2i_resize_crop_u8 per image followed by b2i_predict_u8, on fixture JPEGs and synthetic sizes."""
    from oracle import stb_resize
    imgs = [stb_resize.decode_image(p) for _, p in fixture_paths()]
    rng = np.random.default_rng(5)
    imgs += [rng.integers(0, 256, (h, w, 3), dtype=np.uint8) for h, w in ((224, 224), (90, 301), (640, 480))]
    with pkg.Model(ctx, "resnet18", resnet18_ot, precision=BF16, max_batch=128, micro_batch=32) as m:
        crops = np.stack([ctx.resize_crop(im) for im in imgs])
        # the GPU resize + crop of every fixture is the oracle's, bit for bit
        assert np.array_equal(crops[:NFIX], fixture_crops())
        want = m.predict_u8(crops, k=5, want_logits=True)
        got = m.predict_raw(imgs, k=5, want_logits=True)
        for a, b in zip(got, want):
            assert np.array_equal(a, b)
        again = m.predict_raw(imgs[:3], k=1)           # scratch buffers are reused
        assert np.array_equal(again[0][:, 0], want[0][:3, 0])
        assert m.predict_raw([], k=5)[0].shape == (0, 5)
    # and the crops themselves are the oracle's
    assert np.array_equal(crops[0], stb_resize.resize_crop_224(imgs[0]))


def test_reload_swaps_weights(pkg, ctx, resnet18_ot, tmp_path):
    """b2i_model_reload: after the `train` flow drops a new archive, the model answers with the new
    weights (checked against the oracle on those weights); a bad archive leaves the old ones active."""
    from oracle import ot_format, synth, tch_oracle
    w2 = synth.resnet18_weights(seed=7)
    path2 = str(tmp_path / "resnet18_v2.ot")
    ot_format.write_ot(path2, w2, stem="resnet18")
    img = synth.synthetic_images(3, seed=71)
    with pkg.Model(ctx, "resnet18", resnet18_ot, precision=FP32, max_batch=4) as m:
        _, _, lg1 = m.predict_u8(img, k=5, want_logits=True)
        m.predict_u8(img, k=5)                      # second use: small-batch graph captured on the old buffers
        m.reload(path2)
        _, _, lg2 = m.predict_u8(img, k=5, want_logits=True)
        _, _, ref2 = tch_oracle.predict_u8("resnet18", w2, img, k=5)
        assert np.abs(lg2 - ref2).max() <= tol_for(FP32, ref2)
        assert np.abs(lg2 - lg1).max() > 100 * tol_for(FP32, ref2)
        with pytest.raises(pkg.B2IError):
            m.reload(str(tmp_path / "missing.ot"))
        _, _, lg3 = m.predict_u8(img, k=5, want_logits=True)
        assert np.array_equal(lg3, lg2)


def test_model_load_errors(pkg, ctx, tmp_path, resnet18_ot):
    with pytest.raises(pkg.B2IError) as e:
        pkg.Model(ctx, "vgg16", resnet18_ot)
    assert e.value.code == -1                                   # unknown model: the reference returns None
    with pytest.raises(pkg.B2IError) as e:
        pkg.Model(ctx, "resnet18", str(tmp_path / "nope.ot"))
    assert e.value.code == -2                                   # the reference panics at start-up
    with pytest.raises(pkg.B2IError) as e:
        pkg.Model(ctx, "alexnet", resnet18_ot)                  # wrong archive for the graph
    assert e.value.code == -2 and "missing tensor" in str(e.value)


def test_member_predict_mirrors_reference_semantics(pkg, ctx, resnet18_ot, alexnet_ot, tmp_path):
    """Member::predict (services.rs:475-497) over the fixture directory layout."""
    import shutil
    labels = pkg.load_labels(os.path.join(GOLDEN, "synset_words.txt"))
    assert len(labels) == 1000 and labels[0] == ("n01440764", "tench, Tinca tinca")
    wd = tmp_path / "pretrained_models"
    wd.mkdir()
    os.symlink(resnet18_ot, wd / "resnet18.ot")
    os.symlink(alexnet_ot, wd / "alexnet.ot")
    root = os.path.join(GOLDEN, "imagenet_1k", "train")
    member = pkg.MemberState(ctx, labels, weight_dir=str(wd), image_root=root, max_batch=128, decoder="pil")
    try:
        assert member.predict("vgg16", ["n01491361"]) is None
        ids = list(GV["wnids"])
        res = member.predict("resnet18", ids)
        assert len(res) == NFIX and all(0.0 < p <= 1.0 and isinstance(lbl, str) for p, lbl in res)
        top1 = np.argmax(GV["resnet18_logits"], 1)
        srt = np.sort(GV["resnet18_logits"], 1)
        for i, (p, lbl) in enumerate(res):
            if srt[i, -1] - srt[i, -2] > 2 * bf16_budget(GV["resnet18_logits"]):
                assert lbl == labels[top1[i]][1]
        demo = pkg.sample_predict(labels, ctx=ctx, weight_dir=str(wd), image_path=os.path.join(root, "n01491361", sorted(os.listdir(os.path.join(root, "n01491361")))[0]))
        assert isinstance(demo, str) and demo in dict((l, 1) for _, l in labels)
        res5 = member.predict("alexnet", ids[:2], k=5)
        assert len(res5) == 2 and len(res5[0]) == 5
        with pytest.raises(FileNotFoundError):
            member.predict("resnet18", ["n00000000"])           # missing directory: the reference unwraps
        empty = tmp_path / "imgs" / "n123"
        empty.mkdir(parents=True)
        bad = tmp_path / "imgs" / "n456"
        bad.mkdir()
        (bad / "x.JPEG").write_bytes(b"not a jpeg")
        member.image_root = str(tmp_path / "imgs")
        assert member.predict("resnet18", ["n123"]) == []       # empty dir: id skipped (services.rs:490)
        assert member.predict("resnet18", ["n456"]) is None     # decode failure voids the call (services.rs:492)
        # leader-side accounting on top (services.rs:407-433, 54-81)
        member.image_root = root
        leader = pkg.Leader(member, [(w, dict(labels)[w]) for w in ids])
        pkg.assign_members([0], leader.resnet18, leader.alexnet)
        assert leader.resnet18.assigned_member_ids == [] and leader.alexnet.assigned_member_ids == [0]
        leader.run_job(leader.resnet18, batch=4)                 # nobody assigned: labels skipped, nothing counted
        assert leader.resnet18.finished_prediction_count == 0
        leader.run_job(leader.alexnet, batch=32)
        assert leader.alexnet.finished_prediction_count == NFIX and len(leader.alexnet.query_durations) == NFIX
        rep = pkg.jobs_report(leader.alexnet)
        assert rep["total"] == NFIX and rep["p99_ms"] >= rep["median_ms"]
    finally:
        member.close()


def test_one_bad_image_costs_one_answer_not_the_batch(pkg, ctx, resnet18_ot, tmp_path):
    """b2i_predict_raw_u8_status: a batch of 8 with one undecoded image (NULL) gives 7 answers, bit-identical
    to the same 7 images alone, and one B2I_EDECODE status.  Above the ABI, MemberState.predict_each keeps the
    ids aligned when an empty directory, an undecodable file and a missing directory sit in the middle of a
    batch, and Leader.run_job scores the survivors against their own labels (reference: one id per RPC,
    services.rs:421,485-492)."""
    import importlib
    import shutil
    from oracle import stb_resize
    svc = importlib.import_module("distributed-machine-learning-cluster_b200.service")
    paths = fixture_paths()[:8]
    imgs = [stb_resize.decode_image(p) for _, p in paths]
    with pkg.Model(ctx, "resnet18", resnet18_ot, precision=BF16, max_batch=8) as m:
        holed = list(imgs)
        holed[3] = None
        idx, prob, logits, status = m.predict_raw_status(holed, k=5, want_logits=True)
        assert status.tolist() == [0, 0, 0, pkg._lib.EDECODE, 0, 0, 0, 0]
        assert (idx[3] == -1).all() and (prob[3] == 0).all()
        alone = m.predict_raw(imgs[:3] + imgs[4:], k=5, want_logits=True)
        keep = [0, 1, 2, 4, 5, 6, 7]
        assert np.array_equal(idx[keep], alone[0]) and np.array_equal(logits[keep], alone[2])
        i2, p2, st2 = m.predict_raw_status([None, None], k=1)
        assert st2.tolist() == [pkg._lib.EDECODE] * 2 and (i2 == -1).all()
    labels = pkg.load_labels(os.path.join(GOLDEN, "synset_words.txt"))
    by = dict(labels)
    root = tmp_path / "train"
    ids = []
    for j, (wnid, path) in enumerate(paths):
        (root / wnid).mkdir(parents=True)
        shutil.copy(path, root / wnid / os.path.basename(path))
        ids.append(wnid)
    (root / "n_empty").mkdir()
    (root / "n_bad").mkdir()
    (root / "n_bad" / "x.JPEG").write_bytes(b"\xff\xd8 not really a jpeg")
    ids = ids[:2] + ["n_empty"] + ids[2:4] + ["n_bad"] + ids[4:6] + ["n_gone"] + ids[6:]
    wd = tmp_path / "pretrained_models"
    wd.mkdir()
    os.symlink(resnet18_ot, wd / "resnet18.ot")
    for decoder in ("pil", "auto"):
        member = pkg.MemberState(ctx, labels, weight_dir=str(wd), image_root=str(root), max_batch=16, models=("resnet18",),
                                 decoder=decoder)
        try:
            each = member.predict_each("resnet18", ids)
            assert [e if isinstance(e, str) else "ok" for e in each] == \
                ["ok", "ok", svc.SKIPPED, "ok", "ok", svc.UNDECODABLE, "ok", "ok", svc.MISSING, "ok", "ok"], (decoder, each)
            clean = member.predict_each("resnet18", [w for w, _ in paths])
            if member.decoder == "pil":       # same decoder, same pixels: the answers do not depend on the neighbours
                assert [e for e in each if not isinstance(e, str)] == clean
            assert member.predict("resnet18", ids[:5]) == [e for e in each[:5] if not isinstance(e, str)]
            assert member.predict("resnet18", ids[:6]) is None          # reference semantics of the plain handler
            leader = pkg.Leader(member, [(w, by.get(w, "?")) for w in ids])
            pkg.assign_members([0, 0], leader.resnet18, leader.alexnet)
            leader.alexnet.assigned_member_ids = []
            leader.resnet18.assigned_member_ids = [0]
            lines = []
            leader.run_job(leader.resnet18, batch=11, log=lines.append)
            assert leader.resnet18.finished_prediction_count == 8
            want_correct = sum(1 for (w, _), e in zip(paths, clean) if e[1] == by[w])
            if member.decoder == "pil":
                assert leader.resnet18.correct_prediction_count == want_correct
            assert all(ln.split(" - ")[1].split(":")[0] in by for ln in lines)
        finally:
            member.close()


def test_predict_jpeg_decodes_on_the_gpu(pkg, ctx, resnet18_ot):
    """b2i_predict_jpeg: encoded streams in, top-k out.  Decoders differ by a few LSB, so
    the pixels the GPU decoder produced are read back (b2i_jpeg_decode_u8) and THOSE go through the oracle's
    resize and through b2i_predict_u8: the JPEG path must equal that bit for bit, and sit within the bf16
    budget of the fp32 oracle on the same pixels.  A damaged stream costs its own answer only."""
    from oracle import stb_resize, synth, tch_oracle
    try:
        backend = ctx.jpeg_backend()
    except pkg.B2IError as e:
        if e.code == pkg._lib.ENODEV and "not installed" in str(e):
            pytest.skip("nvJPEG is not installed on this box")
        raise
    print("nvJPEG backends:", backend)
    rows = [ln.rstrip("\n").split("\t") for ln in open(os.path.join(GOLDEN, "fixtures_manifest.tsv"))][1:]
    kind = {r[0]: r[7] for r in rows}
    paths = fixture_paths()
    pick = []
    for want in ("grayscale", "side < 224", "4:2:2", "4:2:0", "widest", "tallest", "square", "ordinary"):
        pick += [p for p in paths if kind[p[0]].startswith(want)][:3]
    blobs = [open(p, "rb").read() for _, p in pick]
    pix = [ctx.jpeg_decode(b) for b in blobs]
    ref = [stb_resize.decode_image(p) for _, p in pick]
    worst = 0
    for a, b, (w, _) in zip(pix, ref, pick):
        assert a.shape == b.shape, (w, a.shape, b.shape)
        d = np.abs(a.astype(int) - b.astype(int))
        worst = max(worst, int(d.max()))
        assert d.mean() < 1.5, (w, kind[w], float(d.mean()))          # same picture, decoder-level differences only
    print("GPU decoder vs libjpeg (PIL) on %d fixtures: max |diff| %d LSB" % (len(pick), worst))
    crops = np.stack([stb_resize.resize_crop_224(p) for p in pix])
    with pkg.Model(ctx, "resnet18", resnet18_ot, precision=BF16, max_batch=32) as m:
        want = m.predict_u8(crops, k=5, want_logits=True)
        idx, prob, logits, status = m.predict_jpeg(blobs, k=5, want_logits=True)
        assert (status == 0).all()
        differ = [i for i in range(len(blobs)) if not np.array_equal(logits[i], want[2][i])]
        assert not differ, ("images whose JPEG-path logits differ from the pixel path", differ, [kind[pick[i][0]] for i in differ],
                            [float(np.abs(logits[i] - want[2][i]).max()) for i in differ])
        assert np.array_equal(idx, want[0]) and np.array_equal(prob, want[1])
        _, _, oracle_logits = tch_oracle.predict_u8("resnet18", synth.resnet18_weights(), crops)
        assert np.abs(logits - oracle_logits).max() <= bf16_budget(oracle_logits)
        # damaged / missing streams in the middle of a batch
        broken = list(blobs[:6])
        broken[1] = b"\xff\xd8\xff\xe0 this is not a jpeg stream"
        broken[4] = b""
        i2, p2, l2, st2 = m.predict_jpeg(broken, k=5, want_logits=True)
        assert st2.tolist() == [0, pkg._lib.EDECODE, 0, 0, pkg._lib.EDECODE, 0]
        keep = [0, 2, 3, 5]
        assert np.array_equal(l2[keep], logits[keep]) and (i2[[1, 4]] == -1).all()
        assert m.predict_jpeg([], k=5)[0].shape == (0, 5)

