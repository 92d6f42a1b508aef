"""The allocation-free resize entry point (b2i_resize_crop_u8_static), for hosts that must not
allocate GPU memory while serving -- e.g. a process whose device memory is handed out once at
start-up.  It has to produce exactly what the allocating entry point produces, call after call,
even though all calls share one fixed scratch array that still holds the previous call's rows.
Collected last on purpose: a failure here cannot hide a failure in the other GPU tests."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def run_static(ctx, img):
    d_in = torch.from_numpy(np.ascontiguousarray(img)).cuda()
    d_out = torch.full((224, 224, 3), 0xAB, dtype=torch.uint8, device="cuda")
    ctx.resize_crop_static(d_in.data_ptr(), img.shape[0], img.shape[1], d_out.data_ptr(),
                           torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    return d_out.cpu().numpy()


def test_static_scratch_matches_allocating_entry_point_on_consecutive_calls(pkg, ctx):
    rng = np.random.default_rng(11)
    # the second and third images are smaller than the first, so rows left over from an earlier
    # call are still sitting in the scratch while they run
    shapes = [(900, 640), (376, 500), (100, 150), (224, 224), (1024, 300)]
    for h, w in shapes:
        img = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
        want = ctx.resize_crop(img)            # existing entry point (allocates its own buffers)
        got = run_static(ctx, img)
        assert np.array_equal(got, want), (h, w)   # integer output: bit-exact, as in the other resize tests
        again = run_static(ctx, img)           # straight away a second time
        assert np.array_equal(again, want), (h, w)


def test_static_scratch_refuses_images_beyond_its_limit(pkg, ctx):
    img = np.zeros((1025, 64, 3), np.uint8)
    d_in = torch.from_numpy(img).cuda()
    d_out = torch.empty((224, 224, 3), dtype=torch.uint8, device="cuda")
    with pytest.raises(pkg.B2IError) as e:
        ctx.resize_crop_static(d_in.data_ptr(), 1025, 64, d_out.data_ptr(), None)
    assert e.value.code == -1 and "1024" in str(e.value)
    # the entry point is still usable afterwards
    ok = np.full((300, 300, 3), 9, np.uint8)
    assert np.array_equal(run_static(ctx, ok), ctx.resize_crop(ok))


def test_static_scratch_refuses_a_capturing_stream(pkg, ctx):
    """The resampling plans are per image and are uploaded from host memory that does not outlive the call,
    so the entry point declines to be recorded into a CUDA graph (B2I_EINVAL) and keeps working afterwards."""
    img = np.full((300, 260, 3), 200, np.uint8)
    d_in = torch.from_numpy(img).cuda()
    d_out = torch.zeros((224, 224, 3), dtype=torch.uint8, device="cuda")
    run_static(ctx, img)                       # the address look-up of the first call happens outside any capture
    side = torch.cuda.Stream()
    graph = torch.cuda.CUDAGraph()
    refused = None
    with torch.cuda.graph(graph, stream=side, capture_error_mode="thread_local"):
        try:
            ctx.resize_crop_static(d_in.data_ptr(), 300, 260, d_out.data_ptr(), torch.cuda.current_stream().cuda_stream)
        except pkg.B2IError as e:
            refused = e
        d_out.add_(0)                          # the capture itself stays valid
    assert refused is not None and refused.code == -1 and "captur" in str(refused)
    torch.cuda.synchronize()
    assert np.array_equal(run_static(ctx, img), ctx.resize_crop(img))
