"""Host half of the resize (target size, crop, resampling plans) computed by the C++ library,
against the numpy restatement -- bit-exact, no GPU needed (b2i_resize_plan / b2i_resize_target)."""
import numpy as np
import pytest

from oracle import stb_resize


@pytest.mark.parametrize("h,w", [(332, 500), (500, 331), (100, 100), (94, 120), (2877, 100), (224, 224), (224, 2304),
                                 (1, 1), (223, 225), (4000, 3000)])
def test_resize_target_matches_oracle(pkg, h, w):
    assert pkg.resize_target(h, w) == stb_resize.resize_target(h, w)


@pytest.mark.parametrize("in_size,out_size,crop", [(500, 297, 36), (150, 336, 56), (224, 224, 0), (300, 300, 38),
                                                   (500, 1191, 483), (2877, 224, 0), (100, 224, 0), (94, 224, 0),
                                                   (2304, 400, 88), (333, 225, 0), (1, 224, 0)])
def test_resampling_plan_is_bit_exact(pkg, in_size, out_size, crop):
    got, n = pkg.resize_plan(in_size, out_size, crop)
    want = stb_resize.plan_axis(in_size, out_size, crop)
    assert n == sum(len(c) for _, c in want)
    for j, ((gi, gc), (wi, wc)) in enumerate(zip(got, want)):
        assert np.array_equal(gi, wi), j
        assert np.array_equal(gc, wc), j          # float32 coefficients, same bits
        assert abs(float(gc.sum(dtype=np.float64)) - 1.0) < 1e-5


def test_plan_rejects_bad_arguments(pkg):
    with pytest.raises(pkg.B2IError):
        pkg.resize_plan(100, 100, 0)             # target below 224
    with pytest.raises(pkg.B2IError):
        pkg.resize_plan(500, 300, 100)           # crop window leaves the image
