"""
fvsrn_extract_color_ex_noalloc (include/fvsrn_b200.h): the colour-extraction pass with its scratch in a static __device__ array
instead of cudaMallocAsync - for callers that cannot allocate (CUDA graph capture, serving loops). No counterpart in the
reference. Checked through the C ABI against the existing entry point fvsrn_extract_color_ex on the same inputs, two calls in a
row with different frames (the shared scratch keeps the previous call's contents: a stale read would show), every channel
mode and output format; plus the size limit. The tolerance is the one tests/test_gpu_color.py uses for this pass (fp32:
2e-6 absolute); min / max are order-independent, so the two entry points are in fact bit-identical.
Collected last on purpose (file name).
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

B, H, W = 2, 96, 160        # 30720 pixels = 7.5 scratch partials: a ragged last block


def _frame(torch, seed, with_nan):
    g = torch.Generator(device="cuda"); g.manual_seed(seed)
    raw = torch.rand((B, 8, H, W), device="cuda", generator=g, dtype=torch.float32)
    raw[:, 4:7] = raw[:, 4:7] * 2 - 1                      # normals in [-1, 1]
    raw[:, 7] = raw[:, 7] * (3.0 + seed) + 0.25 * seed     # depth range differs from call to call
    if with_nan:
        raw[0, 7, 5, 7] = float("nan")                     # an empty pixel: NaN depth propagates through min / max
    return raw.contiguous()


def _call(fn, raw, channel, tonemap, exposure, fmt):
    import torch
    from fvsrn_b200 import _lib
    shape, dtype = {1: ((B, 4, H, W), torch.float32), 2: ((B, H, W), torch.int32), 3: ((B, H, W, 4), torch.float16)}[fmt]
    out = torch.full(shape, -7, dtype=dtype, device="cuda")
    _lib.check(fn(raw.data_ptr(), out.data_ptr(), B, H, W, channel, int(tonemap), float(exposure), fmt,
                  torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()
    return out.cpu().numpy()


def test_noalloc_entry_point_matches_extract_color_ex_on_two_calls_in_a_row(built_lib):
    import torch
    assert torch.cuda.is_available()
    L = built_lib
    first = True
    for seed, with_nan in ((1, False), (2, False), (3, True), (4, False)):     # consecutive calls on different frames
        raw = _frame(torch, seed, with_nan)
        for channel, tonemap in ((2, False), (3, False), (3, True), (0, False), (1, False)):   # depth first: it is the scratch user
            for fmt in (1, 2, 3):
                want = _call(L.fvsrn_extract_color_ex, raw, channel, tonemap, 1.7, fmt)
                got = _call(L.fvsrn_extract_color_ex_noalloc, raw, channel, tonemap, 1.7, fmt)
                again = _call(L.fvsrn_extract_color_ex_noalloc, raw, channel, tonemap, 1.7, fmt)    # second call in a row, same input
                if fmt == 1:
                    np.testing.assert_allclose(got, want, rtol=0, atol=2e-6, equal_nan=True, err_msg=str((seed, channel, fmt)))
                    np.testing.assert_allclose(again, want, rtol=0, atol=2e-6, equal_nan=True, err_msg=str((seed, channel, fmt)))
                else:
                    assert np.array_equal(got.view(np.uint8), want.view(np.uint8)), (seed, channel, fmt)
                    assert np.array_equal(again.view(np.uint8), want.view(np.uint8)), (seed, channel, fmt)
                if channel == 2 and fmt == 1:
                    if with_nan:
                        assert np.isnan(got[:, :3]).all()
                    else:       # a real depth image: normalised to [0, 1] over the whole batch
                        assert abs(float(got[:, :3].min())) <= 1e-6 and abs(float(got[:, :3].max()) - 1.0) <= 1e-6
                        if first:
                            depth = raw[:, 7].cpu().numpy()
                            ref = (depth - depth.min()) / (depth.max() - depth.min())
                            np.testing.assert_allclose(got[:, 0], ref, rtol=0, atol=2e-6)
                            first = False


def test_noalloc_entry_point_rejects_sizes_beyond_its_scratch(built_lib):
    import torch
    from fvsrn_b200 import _lib
    L = built_lib
    raw = _frame(torch, 1, False)
    out = torch.empty((B, 4, H, W), dtype=torch.float32, device="cuda")
    # 2^26 + 8192 pixels > FVSRN_NOALLOC_MAX_PIXELS: refused before any launch (the pointers are never dereferenced)
    status = L.fvsrn_extract_color_ex_noalloc(raw.data_ptr(), out.data_ptr(), 1, 8193, 8192, 2, 0, 1.0, 1, None)
    assert status != 0
    msg = L.fvsrn_last_error().decode()
    assert "FVSRN_NOALLOC_MAX_PIXELS" in msg and "exceeds" in msg
    with pytest.raises(RuntimeError, match="null device pointer"):
        _lib.check(L.fvsrn_extract_color_ex_noalloc(None, out.data_ptr(), B, H, W, 2, 0, 1.0, 1, None))
    # an empty batch is not an error (as with fvsrn_extract_color_ex)
    assert L.fvsrn_extract_color_ex_noalloc(raw.data_ptr(), out.data_ptr(), 0, 8192, 8192, 2, 0, 1.0, 1, None) == 0
