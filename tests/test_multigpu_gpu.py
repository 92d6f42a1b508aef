"""In-process sharding over several GPUs of one box (b2i_predict_u8 / b2i_predict_u8_on): the
batch is split contiguously, every GPU runs its slice on its own streams, only top-k comes back;
results equal the single-GPU run image for image.  Needs >= 2 GPUs (`gpurun --gpus 2`); skipped on
the single-GPU test box."""
import threading

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

BF16 = 0


@pytest.fixture(scope="module")
def two_gpus(pkg):
    if pkg.device_count() < 2:
        pytest.skip("needs two GPUs")
    return pkg.device_count()


def test_sharded_predict_equals_single_gpu(pkg, ctx, two_gpus, resnet18_ot):
    from oracle import synth
    imgs = synth.synthetic_images(37, seed=31)                     # odd: ragged shards (19 + 18)
    with pkg.Model(ctx, "resnet18", resnet18_ot, precision=BF16, max_batch=40) as m1:
        i1, p1, l1 = m1.predict_u8(imgs, k=5, want_logits=True)
    with pkg.Context((0, 1)) as c2, pkg.Model(c2, "resnet18", resnet18_ot, precision=BF16, max_batch=20) as m2:
        assert m2.num_devices == 2
        i2, p2, l2 = m2.predict_u8(imgs, k=5, want_logits=True)    # 37 <= 2 x 20
        assert np.array_equal(i2, i1) and np.array_equal(l2, l1) and np.array_equal(p2, p1)
        # one image, fewer images than GPUs, empty batch
        ia, pa = m2.predict_u8(imgs[:1], k=5)
        assert np.array_equal(ia, i1[:1])
        i0, _ = m2.predict_u8(imgs[:0], k=5)
        assert i0.shape == (0, 5)
        with pytest.raises(pkg.B2IError):
            m2.predict_u8(synth.synthetic_images(41, seed=1), k=5)  # > 2 x max_batch
        # a job restricted to one GPU of the pair (the dual-job split, services.rs:199-211)
        ib, pb = m2.predict_u8(imgs[:20], k=5, devices=(1, 1))
        assert np.array_equal(ib, i1[:20])
        # asynchronous pair on both GPUs
        oi, op = np.empty((37, 5), np.int32), np.empty((37, 5), np.float32)
        m2.wait(m2.submit_u8(imgs, oi, op))
        assert np.array_equal(oi, i1)


def test_two_jobs_run_concurrently_on_disjoint_gpus(pkg, two_gpus, resnet18_ot, alexnet_ot):
    """ResNet-18 on GPU 0 and AlexNet on GPU 1 at the same time (Leader::predict joins both jobs,
    services.rs:146-151): answers equal the ones each model gives alone."""
    from oracle import synth
    imgs = synth.synthetic_images(16, seed=33)
    with pkg.Context((0, 1)) as c2, \
            pkg.Model(c2, "resnet18", resnet18_ot, precision=BF16, max_batch=16) as mr, \
            pkg.Model(c2, "alexnet", alexnet_ot, precision=BF16, max_batch=16) as ma:
        want_r = mr.predict_u8(imgs, k=5, devices=(0, 1))
        want_a = ma.predict_u8(imgs, k=5, devices=(1, 1))
        got = {}

        def job(name, model, dev):
            for _ in range(5):
                got[name] = model.predict_u8(imgs, k=5, devices=(dev, 1))

        th = [threading.Thread(target=job, args=("r", mr, 0)), threading.Thread(target=job, args=("a", ma, 1))]
        for t in th:
            t.start()
        for t in th:
            t.join()
        assert np.array_equal(got["r"][0], want_r[0]) and np.array_equal(got["a"][0], want_a[0])
