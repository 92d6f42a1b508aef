"""Per-kernel parity on the GPU, through the C ABI's single-kernel entry points, against torch CPU
fp32 on the same (bf16-rounded, for the bf16 path) inputs.

Tolerances: the bf16 path rounds its output once to bf16 (relative 2^-8) after fp32 accumulation, so
|got - ref| <= 2^-7 * |ref| + 1e-3 * scale; the fp32 path must agree to 1e-4 relative to the
output scale (BASELINE.json: 1e-4 max-abs on O(1) logits)."""
import os

import numpy as np
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

BF16, FP32, FP16 = 0, 1, 2
GATHER, IM2COL = 0, 1


def bf16_round(a):
    return torch.as_tensor(np.asarray(a, dtype=np.float32)).to(torch.bfloat16).to(torch.float32)


def fp16_round(a):
    return torch.as_tensor(np.asarray(a, dtype=np.float32)).to(torch.float16).to(torch.float32)


def round16(a, precision):
    return fp16_round(a) if precision == FP16 else bf16_round(a)


def rnd(shape, seed, scale=1.0):
    return (torch.randn(*shape, generator=torch.Generator().manual_seed(seed)) * scale)


def report(got, ref, tol_rel, tol_abs, what):
    got = np.asarray(got, dtype=np.float64)
    ref = np.asarray(ref, dtype=np.float64)
    assert got.shape == ref.shape, (what, got.shape, ref.shape)
    err = np.abs(got - ref)
    bound = tol_rel * np.abs(ref) + tol_abs
    bad = ~(err <= bound)  # NaN counts as bad
    if bad.any():
        flat = bad.reshape(-1, bad.shape[-1])
        rows = np.nonzero(flat.any(1))[0]
        cols = np.nonzero(flat.any(0))[0]
        msg = ("%s: %d / %d outside tolerance; max err %.4g (ref scale %.4g); nan %d; bad rows %d (first %s, mod128 %s); "
               "bad cols %d (first %s)" % (what, bad.sum(), bad.size, np.nanmax(err), np.abs(ref).max(), np.isnan(got).sum(),
                                           len(rows), rows[:8].tolist(), sorted(set((rows % 128).tolist()))[:16],
                                           len(cols), cols[:8].tolist()))
        raise AssertionError(msg)


def conv_case(ctx, N, H, W, C, Cout, k, stride, pad, *, precision, conv_mode, weights_tma=1, bias=False, bn=False,
              residual=False, relu=False, tile_n=0, seed=0):
    x = rnd((N, H, W, C), seed)
    w = rnd((Cout, C, k, k), seed + 1, (2.0 / (C * k * k)) ** 0.5)
    b = rnd((Cout,), seed + 2, 0.5) if bias else None
    bnp = None
    if bn:
        g = torch.Generator().manual_seed(seed + 3)
        bnp = (torch.rand(Cout, generator=g) + 0.5, torch.randn(Cout, generator=g) * 0.1,
               torch.randn(Cout, generator=g) * 0.1, torch.rand(Cout, generator=g) + 0.5)
    OH, OW = (H + 2 * pad - k) // stride + 1, (W + 2 * pad - k) // stride + 1
    res = rnd((N, OH, OW, Cout), seed + 4) if residual else None
    if precision != FP32:
        x, res = round16(x, precision), (round16(res, precision) if residual else None)
    got = ctx.conv2d(x.numpy(), w.numpy(), bias=None if b is None else b.numpy(),
                     bn=None if bnp is None else [t.numpy() for t in bnp], stride=stride, pad=pad,
                     residual=None if res is None else res.numpy(), relu=relu, precision=precision,
                     conv_mode=conv_mode, weights_tma=weights_tma, tile_n=tile_n)
    # reference: fold exactly as the library does, in fp32, then round weights for the bf16 path
    wf, shift = w, (b if b is not None else torch.zeros(Cout))
    if bnp is not None:
        gamma, beta, mean, var = bnp
        inv = 1.0 / torch.sqrt(var.double() + 1e-5)
        wf = w * (gamma.double() * inv).float().view(-1, 1, 1, 1)
        shift = (beta.double() - mean.double() * gamma.double() * inv + shift.double() * gamma.double() * inv * (1 if b is not None else 0)).float()
    if precision != FP32:
        wf = round16(wf, precision)
    ref = F.conv2d(x.permute(0, 3, 1, 2).double(), wf.double(), shift.double(), stride, pad).permute(0, 2, 3, 1)
    if res is not None:
        ref = ref + res.double()
    if relu:
        ref = ref.clamp_min(0)
    scale = float(ref.abs().max())
    if precision == BF16:
        report(got, ref.numpy(), 2.0 ** -7, 1e-3 * scale, "conv bf16")
    elif precision == FP16:   # one rounding to half (relative 2^-11) after fp32 accumulation
        report(got, ref.numpy(), 2.0 ** -10, 2e-4 * scale, "conv fp16")
    else:
        report(got, ref.numpy(), 0.0, 1e-5 * scale, "conv fp32")


# ---- tcgen05 path: every A-operand mode on the same geometries ------------------------------
IM2COL_ONE_TILE = 2
MODES = [("gather_cpasyncB", GATHER, 0), ("gather_tmaB", GATHER, 1), ("im2col", IM2COL, 1),
         ("im2col_onetile", IM2COL_ONE_TILE, 1), ("pair_im2col", 5, 1)]


@pytest.mark.parametrize("name,mode,btma", MODES, ids=[m[0] for m in MODES])
def test_conv3x3_c64_partial_tile(ctx, name, mode, btma):
    conv_case(ctx, 2, 12, 12, 64, 64, 3, 1, 1, precision=BF16, conv_mode=mode, weights_tma=btma)


@pytest.mark.parametrize("name,mode,btma", MODES, ids=[m[0] for m in MODES])
def test_conv3x3_resnet_layer1_block(ctx, name, mode, btma):
    conv_case(ctx, 3, 56, 56, 64, 64, 3, 1, 1, precision=BF16, conv_mode=mode, weights_tma=btma, bn=True, residual=True,
              relu=True, seed=10)


@pytest.mark.parametrize("name,mode,btma", MODES, ids=[m[0] for m in MODES])
def test_conv3x3_stride2_c128_to_256(ctx, name, mode, btma):
    conv_case(ctx, 3, 28, 28, 128, 256, 3, 2, 1, precision=BF16, conv_mode=mode, weights_tma=btma, bn=True, relu=True, seed=20)


@pytest.mark.parametrize("name,mode,btma", MODES, ids=[m[0] for m in MODES])
def test_conv1x1_stride2_downsample(ctx, name, mode, btma):
    conv_case(ctx, 5, 14, 14, 256, 512, 1, 2, 0, precision=BF16, conv_mode=mode, weights_tma=btma, bn=True, seed=30)


@pytest.mark.parametrize("name,mode,btma", MODES, ids=[m[0] for m in MODES])
def test_conv3x3_tiny_maps_straddle_images(ctx, name, mode, btma):
    """7x7 maps: one 128-row tile covers 2.6 images (layer4 geometry)."""
    conv_case(ctx, 9, 7, 7, 512, 512, 3, 1, 1, precision=BF16, conv_mode=mode, weights_tma=btma, bn=True, residual=True,
              relu=True, seed=40)


@pytest.mark.parametrize("name,mode,btma", MODES, ids=[m[0] for m in MODES])
def test_conv5x5_alexnet_both_geometries(ctx, name, mode, btma):
    conv_case(ctx, 2, 27, 27, 64, 192, 5, 2, 1, precision=BF16, conv_mode=mode, weights_tma=btma, bias=True, relu=True, seed=50)
    conv_case(ctx, 2, 27, 27, 64, 192, 5, 1, 2, precision=BF16, conv_mode=mode, weights_tma=btma, bias=True, relu=True, seed=51)


@pytest.mark.parametrize("name,mode,btma", MODES, ids=[m[0] for m in MODES])
@pytest.mark.parametrize("tile_n", [32, 64, 128, 256])
def test_conv_tile_n_sweep(ctx, name, mode, btma, tile_n):
    conv_case(ctx, 2, 14, 14, 128, 256, 3, 1, 1, precision=BF16, conv_mode=mode, weights_tma=btma, tile_n=tile_n, seed=60)


def test_conv3x3_c64_halo_patch_kernel(ctx, variant=3):
    """ResNet layer1 geometry through the halo-patch kernel: taps as shifted descriptor views."""
    conv_case(ctx, 3, 56, 56, 64, 64, 3, 1, 1, precision=BF16, conv_mode=variant, bn=True, residual=True, relu=True, seed=200)
    conv_case(ctx, 2, 24, 40, 64, 64, 3, 1, 1, precision=BF16, conv_mode=variant, bn=True, seed=201)      # H not a multiple of 16
    conv_case(ctx, 5, 7, 8, 64, 64, 3, 1, 1, precision=BF16, conv_mode=variant, bias=True, relu=True, seed=202)


def test_conv3x3_flattened_rows_kernel(ctx):
    """3x3 / s1 / p1 through the flattened padded-row kernel (conv_mode 6): nine shifted views of one
    row patch, CTA pair, resident weights.  Row pitch 32 (W = 28), 16 (W <= 14) and 64 (W <= 62); tiles
    that straddle images, a last tile that runs past the batch, odd sizes, residual / ReLU / plain bias."""
    conv_case(ctx, 3, 28, 28, 128, 128, 3, 1, 1, precision=BF16, conv_mode=6, bn=True, residual=True, relu=True, seed=300)
    conv_case(ctx, 37, 28, 28, 128, 128, 3, 1, 1, precision=BF16, conv_mode=6, bn=True, relu=True, seed=301)   # several tiles per CTA pair
    conv_case(ctx, 5, 14, 14, 128, 128, 3, 1, 1, precision=BF16, conv_mode=6, bn=True, residual=True, seed=302)   # pitch 16
    conv_case(ctx, 2, 9, 13, 64, 64, 3, 1, 1, precision=BF16, conv_mode=6, bias=True, relu=True, seed=303)       # H != W, one slice
    conv_case(ctx, 1, 56, 56, 64, 64, 3, 1, 1, precision=BF16, conv_mode=6, bn=True, residual=True, relu=True, seed=304)   # pitch 64
    conv_case(ctx, 2, 30, 30, 64, 128, 3, 1, 1, precision=BF16, conv_mode=6, bias=True, seed=305)                # W + 2 == pitch


def test_conv_pair_multi_subtile(ctx):
    """CTA-pair im2col kernel with two 128-row sub-tiles per CTA (conv_mode 8): one 256-pixel im2col request
    per k-block, one or two accumulator sets, tiles that end inside or past the batch."""
    conv_case(ctx, 9, 7, 7, 512, 512, 3, 1, 1, precision=BF16, conv_mode=8, tile_n=256, bn=True, residual=True, relu=True, seed=320)
    conv_case(ctx, 37, 7, 7, 512, 512, 3, 1, 1, precision=BF16, conv_mode=8, tile_n=128, bn=True, residual=True, relu=True, seed=321)
    conv_case(ctx, 3, 56, 56, 64, 128, 3, 2, 1, precision=BF16, conv_mode=8, tile_n=128, bn=True, relu=True, seed=322)
    conv_case(ctx, 2, 14, 14, 256, 256, 3, 1, 1, precision=BF16, conv_mode=8, tile_n=128, bias=True, seed=323)


def test_conv_fp16_storage_mode(ctx):
    """PREC_FP16: the same tcgen05 kernels with IEEE-half activations / weights (instruction descriptor formats,
    epilogue pack / unpack, residual): every kernel family once, against fp32 math on half-rounded inputs."""
    conv_case(ctx, 3, 56, 56, 64, 64, 3, 1, 1, precision=FP16, conv_mode=3, bn=True, residual=True, relu=True, seed=400)      # halo patch
    conv_case(ctx, 5, 28, 28, 128, 128, 3, 1, 1, precision=FP16, conv_mode=6, bn=True, residual=True, relu=True, seed=401)   # flattened rows
    conv_case(ctx, 3, 56, 56, 64, 128, 3, 2, 1, precision=FP16, conv_mode=7, bn=True, relu=True, seed=402)                  # pair, resident weights
    conv_case(ctx, 9, 7, 7, 512, 512, 3, 1, 1, precision=FP16, conv_mode=8, tile_n=256, bn=True, residual=True, relu=True, seed=403)   # pair, two sub-tiles
    conv_case(ctx, 3, 14, 14, 256, 256, 3, 1, 1, precision=FP16, conv_mode=IM2COL, bias=True, seed=404)                     # persistent im2col
    conv_case(ctx, 2, 12, 12, 64, 64, 3, 1, 1, precision=FP16, conv_mode=GATHER, bn=True, relu=True, seed=405)              # one-tile gather
    conv_case(ctx, 2, 67, 67, 3, 64, 11, 4, 2, precision=FP16, conv_mode=GATHER, bias=True, relu=True, seed=406)            # small-channel gather


def test_conv_pair_resident_weights(ctx):
    """CTA-pair im2col kernel keeping all weight k-blocks resident (conv_mode 7): the stride-2 first conv of
    ResNet layer2, a 1x1, and a layer whose weights do not fit (falls back to streaming them)."""
    conv_case(ctx, 3, 56, 56, 64, 128, 3, 2, 1, precision=BF16, conv_mode=7, bn=True, relu=True, seed=310)
    conv_case(ctx, 9, 28, 28, 128, 128, 3, 1, 1, precision=BF16, conv_mode=7, bn=True, residual=True, relu=True, seed=311)
    conv_case(ctx, 5, 56, 56, 64, 128, 1, 2, 0, precision=BF16, conv_mode=7, bn=True, seed=312)
    conv_case(ctx, 3, 14, 14, 256, 256, 3, 1, 1, precision=BF16, conv_mode=7, tile_n=256, bn=True, residual=True, relu=True, seed=313)


@pytest.mark.parametrize("btma", [0, 1], ids=["smallc_cpasyncB", "smallc_tmaB"])
def test_conv_first_layers_rgb(ctx, btma):
    """C = 3 layers: ResNet conv1 (7x7 s2 p3) and AlexNet features.0 (11x11 s4 p2)."""
    conv_case(ctx, 2, 64, 64, 3, 64, 7, 2, 3, precision=BF16, conv_mode=GATHER, weights_tma=btma, bn=True, relu=True, seed=70)
    conv_case(ctx, 2, 67, 67, 3, 64, 11, 4, 2, precision=BF16, conv_mode=GATHER, weights_tma=btma, bias=True, relu=True, seed=71)


def test_linear_tiled_bf16(ctx):
    for (M, K, Cout, seed) in [(5, 512, 1000, 80), (300, 1024, 4096, 81), (1, 9216, 64, 82)]:
        x = bf16_round(rnd((M, K), seed))
        w = rnd((Cout, K), seed + 1, (1.0 / K) ** 0.5)
        b = rnd((Cout,), seed + 2, 0.1)
        got = ctx.linear(x.numpy(), w.numpy(), b.numpy(), relu=False, precision=BF16)
        ref = (x.double() @ bf16_round(w).double().T + b.double()).numpy()
        report(got, ref, 0.0, 2e-5 * np.abs(ref).max() + 1e-6, "linear bf16 (fp32 out)")


def test_linear_tiled_fp16(ctx):
    for (M, K, Cout, seed) in [(5, 512, 1000, 83), (300, 1024, 4096, 84)]:
        x = fp16_round(rnd((M, K), seed))
        w = rnd((Cout, K), seed + 1, (1.0 / K) ** 0.5)
        b = rnd((Cout,), seed + 2, 0.1)
        got = ctx.linear(x.numpy(), w.numpy(), b.numpy(), relu=False, precision=FP16)
        ref = (x.double() @ fp16_round(w).double().T + b.double()).numpy()
        report(got, ref, 0.0, 2e-5 * np.abs(ref).max() + 1e-6, "linear fp16 (fp32 out)")


@pytest.mark.parametrize("precision", [BF16, FP16], ids=["bf16", "fp16"])
def test_linear_split_k(ctx, precision):
    """Split-K FC (the k-blocks of one output tile spread over several CTAs, partial sums added in split order):
    same answer as the one-CTA-per-tile kernel up to fp32 summation order, identical from run to run, for ragged
    splits (k-blocks not divisible), more splits than k-blocks, row counts on both sides of a 128-row tile, ReLU."""
    rd = fp16_round if precision == FP16 else bf16_round
    for (M, K, Cout, splits, relu, seed) in [(32, 9216, 4096, 4, True, 120), (5, 4096, 1000, 8, False, 121), (130, 1024, 192, 3, True, 122),
                                             (256, 512, 1000, 2, False, 123), (3, 320, 64, 16, False, 124), (1, 9216, 4096, 7, True, 125)]:
        tile_n = 128 if Cout % 128 == 0 and seed != 122 else 64
        x = rd(rnd((M, K), seed))
        w = rnd((Cout, K), seed + 1, (1.0 / K) ** 0.5)
        b = rnd((Cout,), seed + 2, 0.1)
        got = ctx.linear(x.numpy(), w.numpy(), b.numpy(), relu=relu, precision=precision, tile_n=tile_n, splits=splits)
        ref = x.double() @ rd(w).double().T + b.double()
        ref = (ref.clamp_min(0) if relu else ref).numpy()
        report(got, ref, 0.0, 2e-5 * np.abs(ref).max() + 1e-6, "linear split-K x%d (fp32 out)" % splits)
        again = ctx.linear(x.numpy(), w.numpy(), b.numpy(), relu=relu, precision=precision, tile_n=tile_n, splits=splits)
        assert np.array_equal(got, again)                       # ordered sum: reproducible
        one = ctx.linear(x.numpy(), w.numpy(), b.numpy(), relu=relu, precision=precision, tile_n=64)
        assert np.abs(got - one).max() <= 2e-5 * np.abs(ref).max() + 1e-7


def test_conv_and_linear_fp32_mode(ctx):
    conv_case(ctx, 2, 20, 20, 64, 128, 3, 1, 1, precision=FP32, conv_mode=GATHER, bn=True, residual=True, relu=True, seed=90)
    conv_case(ctx, 2, 33, 33, 3, 64, 7, 2, 3, precision=FP32, conv_mode=GATHER, bn=True, relu=True, seed=91)
    conv_case(ctx, 3, 14, 14, 256, 512, 1, 2, 0, precision=FP32, conv_mode=GATHER, bn=True, seed=92)
    x, w, b = rnd((7, 4096), 93), rnd((1000, 4096), 94, 0.02), rnd((1000,), 95, 0.1)
    got = ctx.linear(x.numpy(), w.numpy(), b.numpy(), precision=FP32)
    ref = (x.double() @ w.double().T + b.double()).numpy()
    report(got, ref, 0.0, 1e-5 * np.abs(ref).max(), "linear fp32")


# ---- bandwidth kernels --------------------------------------------------------------------
@pytest.mark.parametrize("precision", [BF16, FP32, FP16], ids=["bf16", "fp32", "fp16"])
def test_maxpool(ctx, precision):
    for (N, H, W, C, k, s, p, seed) in [(2, 112, 112, 64, 3, 2, 1, 100), (3, 55, 55, 64, 3, 2, 0, 101), (2, 13, 13, 192, 3, 2, 0, 102),
                                        (1, 6, 6, 256, 3, 2, 0, 103)]:
        x = rnd((N, H, W, C), seed)
        if precision != FP32:
            x = round16(x, precision)
        got = ctx.maxpool(x.numpy(), k, s, p, precision=precision)
        ref = F.max_pool2d(x.permute(0, 3, 1, 2), k, s, p).permute(0, 2, 3, 1).numpy()
        assert np.array_equal(got, ref)  # max of representable values: exact in both dtypes


@pytest.mark.parametrize("precision", [BF16, FP32, FP16], ids=["bf16", "fp32", "fp16"])
def test_adaptive_avgpool(ctx, precision):
    for (N, H, W, C, oh, ow, seed) in [(4, 7, 7, 512, 1, 1, 110), (3, 2, 2, 256, 6, 6, 111), (2, 6, 6, 256, 6, 6, 112), (2, 13, 13, 256, 6, 6, 113)]:
        x = rnd((N, H, W, C), seed)
        if precision != FP32:
            x = round16(x, precision)
        got = ctx.adaptive_avgpool(x.numpy(), oh, ow, precision=precision)
        ref = F.adaptive_avg_pool2d(x.permute(0, 3, 1, 2).double(), (oh, ow)).permute(0, 2, 3, 1).numpy()
        if precision == BF16:
            report(got, ref, 2.0 ** -8, 1e-6, "avgpool bf16")
        elif precision == FP16:
            report(got, ref, 2.0 ** -11, 1e-6, "avgpool fp16")
        else:
            report(got, ref, 0.0, 2e-6, "avgpool fp32")


def test_normalize(ctx):
    from oracle import synth, tch_oracle
    img = synth.synthetic_images(3, seed=9)
    ref = tch_oracle.normalize_u8(img).permute(0, 2, 3, 1).numpy()
    got = ctx.normalize(img, precision=FP32)
    assert np.array_equal(got[..., 3], np.zeros_like(got[..., 3]))
    assert np.abs(got[..., :3] - ref).max() <= 2.4e-7 * 3  # same IEEE ops; torch may fuse the broadcast differently by 1 ulp
    gotb = ctx.normalize(img, precision=BF16)
    assert np.array_equal(gotb[..., :3], bf16_round(got[..., :3]).numpy())
    goth = ctx.normalize(img, precision=FP16)
    assert np.array_equal(goth[..., :3], fp16_round(got[..., :3]).numpy())


def test_softmax_topk(ctx):
    from oracle import tch_oracle
    lg = rnd((37, 1000), 120, 3.0)
    lg[3, 10] = lg[3, 500] = lg[3].max() + 1.0  # exact tie: lower index first
    for k in (1, 5):
        idx, prob = ctx.softmax_topk(lg.numpy(), k)
        ridx, rprob = tch_oracle.softmax_topk(lg, k)
        assert np.array_equal(idx[np.arange(37) != 3], ridx.numpy()[np.arange(37) != 3])
        assert np.abs(prob - rprob.numpy()).max() <= 1e-6
    assert idx[3, 0] == 10 and idx[3, 1] == 500
    small = rnd((2, 10), 121)
    idx, prob = ctx.softmax_topk(small.numpy(), 3)
    ridx, rprob = tch_oracle.softmax_topk(small, 3)
    assert np.array_equal(idx, ridx.numpy()) and np.abs(prob - rprob.numpy()).max() <= 1e-6


def test_resize_crop_bit_exact_vs_oracle(ctx):
    """Integer output: bit-exact against the stb restatement, on up-/down-sampling, square and
    extreme-aspect inputs, and on the committed fixture images."""
    from conftest import GOLDEN
    from oracle import stb_resize
    rng = np.random.default_rng(7)
    for (h, w) in [(376, 500), (100, 150), (224, 224), (300, 224), (94, 500), (224, 400), (333, 333), (1100, 260)]:
        img = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
        assert np.array_equal(ctx.resize_crop(img), stb_resize.resize_crop_224(img)), (h, w)
    gv = np.load(os.path.join(GOLDEN, "golden_vectors.npz"))
    root = os.path.join(GOLDEN, "imagenet_1k", "train")
    for wnid in ("n03825788", "n02840245"):
        p = os.path.join(root, wnid, sorted(os.listdir(os.path.join(root, wnid)))[0])
        assert np.array_equal(ctx.resize_crop(stb_resize.decode_image(p)), gv["crop_" + wnid]), wnid
