"""Parity properties at BASELINE.json's full sizes (batch 256 and 1024 per GPU), where running the
CPU oracle on every image would take minutes: properties that hold for the reference operator by
construction (it processes images independently, one by one, services.rs:484-495) and so must hold
bit-exactly for the batched GPU path:

  * an image's logits do not depend on the batch it travels in (split / permutation invariance),
  * repeated calls are idempotent,
  * top-k probabilities are sorted, in (0, 1], and sum to at most 1,
and a spot check of 4 images out of the 1024 against the oracle."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

BF16, FP32, FP16 = 0, 1, 2


@pytest.mark.parametrize("precision", [BF16, FP16], ids=["bf16", "fp16"])
@pytest.mark.parametrize("model_name", ["resnet18", "alexnet"])
def test_full_batch_properties_16bit(pkg, ctx, resnet18_ot, alexnet_ot, model_name, precision):
    from oracle import synth, tch_oracle
    ot = resnet18_ot if model_name == "resnet18" else alexnet_ot
    n = 1024
    base = synth.synthetic_images(256, seed=51)
    imgs = np.concatenate([base, base[::-1], np.roll(base, 7, axis=0), 255 - base])   # 1024 images, many repeats
    with pkg.Model(ctx, model_name, ot, precision=precision, max_batch=n, micro_batch=256) as m:
        idx, prob, lg = m.predict_u8(imgs, k=5, want_logits=True)
        # idempotence
        idx2, prob2, lg2 = m.predict_u8(imgs, k=5, want_logits=True)
        assert np.array_equal(lg, lg2) and np.array_equal(idx, idx2) and np.array_equal(prob, prob2)
        # the same image gives the same logits wherever it sits in the batch
        assert np.array_equal(lg[256:512], lg[:256][::-1])
        assert np.array_equal(lg[512:768], np.roll(lg[:256], 7, axis=0))
        # split invariance: one call of 1024 == calls of 256, 300, 468
        parts = [m.predict_u8(imgs[a:b], k=5, want_logits=True)[2] for a, b in ((0, 256), (256, 556), (556, 1024))]
        assert np.array_equal(np.concatenate(parts), lg)
        # a random permutation permutes the results
        perm = np.random.default_rng(0).permutation(n)
        _, _, lgp = m.predict_u8(imgs[perm], k=5, want_logits=True)
        assert np.array_equal(lgp, lg[perm])
    # probabilities: sorted, in (0, 1], top-5 mass <= 1, consistent with the logits
    assert (prob[:, :-1] >= prob[:, 1:]).all() and (prob > 0).all() and (prob <= 1).all()
    assert (prob.sum(1) <= 1 + 1e-5).all()
    assert np.array_equal(idx[:, 0], lg.argmax(1))
    # spot check against the oracle
    pick = [0, 255, 600, 1023]
    _, _, ref = tch_oracle.predict_u8(model_name, synth.weights(model_name), imgs[pick], k=5)
    # bf16: the dtype's own distance from the fp32 oracle (see test_model_gpu.py); fp16: the flat 2e-2 bar
    assert np.abs(lg[pick] - ref).max() <= (3e-2 if precision == BF16 else 2e-2) * max(1.0, float(ref.std()))


def test_full_batch_fp32_mode_matches_oracle_sample(pkg, ctx, resnet18_ot):
    from oracle import synth, tch_oracle
    imgs = synth.synthetic_images(256, seed=52)
    with pkg.Model(ctx, "resnet18", resnet18_ot, precision=FP32, max_batch=256) as m:
        idx, prob, lg = m.predict_u8(imgs, k=5, want_logits=True)
    pick = [0, 100, 255]
    ridx, rprob, ref = tch_oracle.predict_u8("resnet18", synth.resnet18_weights(), imgs[pick], k=5)
    assert np.abs(lg[pick] - ref).max() <= 1e-4 * max(1.0, float(ref.std()))
    assert np.array_equal(idx[pick], ridx)
    assert np.abs(prob[pick] - rprob).max() <= 1e-5
