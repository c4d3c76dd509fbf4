"""nsof_ssim_u8_noalloc (include/nsof.h): the allocation-free variant of the SSIM entry point -- its per-CTA scratch is a
file-scope `static __device__` array, for callers that may not call cudaMalloc (CUDA graph capture, serving loops).  No
counterpart in the reference.  Called through the C ABI and compared with
the existing entry point nsof_ssim_u8 on two calls in a row (the scratch keeps the previous call's contents).  This
file sorts last on purpose: a failure here must not hide other tests."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_ssim_noalloc_matches_the_allocating_entry_point(nsof):
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from oracle import frames_oracle as fo
    eng = nsof.Engine(0)   # own handle: the test does not depend on what other tests left behind
    lib, h = eng.lib, eng.handle
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    rng = np.random.default_rng(11)
    out_a = torch.zeros(1, dtype=torch.float64, device="cuda")
    out_b = torch.zeros(1, dtype=torch.float64, device="cuda")
    # first call: a 1080p BGR frame pair through the channel-2 view (pixel stride 3), 2040 tiles; second call: a smaller
    # grey pair (fewer tiles: the tail of the shared scratch still holds the first call's partial sums)
    for shape, stride in (((1080, 1920, 3), 3), ((301, 457), 1)):
        true = rng.integers(0, 256, shape, dtype=np.uint8)
        pred = np.clip(true.astype(np.int16) + rng.integers(-25, 26, shape), 0, 255).astype(np.uint8)
        da, db = torch.from_numpy(true).cuda(), torch.from_numpy(pred).cuda()
        H, W = shape[:2]
        off = 2 if stride == 3 else 0
        args = (h, da.data_ptr() + off, W * stride, stride, db.data_ptr() + off, W * stride, stride, H, W, 255.0)
        lib.check(lib.nsof_ssim_u8(*args, out_a.data_ptr(), st), "nsof_ssim_u8")
        before = torch.cuda.memory_allocated()
        lib.check(lib.nsof_ssim_u8_noalloc(*args, out_b.data_ptr(), st), "nsof_ssim_u8_noalloc")
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated() == before
        a, b = float(out_a.item()), float(out_b.item())
        assert a == b, (shape, a, b)   # same kernels, same order of operations: identical bits
        ref = fo.ssim_u8(true[..., 2] if stride == 3 else true, pred[..., 2] if stride == 3 else pred)
        assert abs(b - ref) <= 1e-12, (shape, b, ref)   # the tolerance of test_ssim_vs_oracle
    # beyond the static scratch: 4096 tiles of 32 x 32 interior pixels
    big = torch.zeros((2100, 2100), dtype=torch.uint8, device="cuda")
    with pytest.raises(NotImplementedError):
        lib.check(lib.nsof_ssim_u8_noalloc(h, big.data_ptr(), 2100, 1, big.data_ptr(), 2100, 1, 2100, 2100, 255.0,
                                           out_b.data_ptr(), st), "nsof_ssim_u8_noalloc")
    lib.check(lib.nsof_ssim_u8_noalloc(h, big.data_ptr(), 2100, 1, big.data_ptr(), 2100, 1, 2054, 2054, 255.0,
                                       out_b.data_ptr(), st), "nsof_ssim_u8_noalloc")   # exactly 64 x 64 tiles
    torch.cuda.synchronize()
    assert float(out_b.item()) == 1.0
    eng.close()
