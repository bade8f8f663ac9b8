"""
Colour extraction: IImageEvaluator::ExtractColor (renderer/iimage_evaluator.cpp:26-135 with the kernels of
renderer/iimage_evaluator_cuda.cu:21-80, 145-260) as a pass over the raw tensor (fvsrn_extract_color_ex) and FUSED into
the DVR kernel's pixel write (fvsrn_dvr_params.output_format), against a NumPy restatement of the reference's kernels:
channel table, rgbaToInt packing (renderer_utils.cuh:48-57), ACES tonemapping + gamma.
"""
import ctypes as C

import numpy as np
import pytest

from conftest import Golden
from test_gpu_parity import env, _volume, _evaluator  # noqa: F401

pytestmark = pytest.mark.gpu


def ref_tonemap(rgb, max_exposure):
    """tonemappingFunction, iimage_evaluator_cuda.cu:145-166 (float32)."""
    v = (rgb / np.float32(max_exposure)).astype(np.float32)
    v = (v * (np.float32(2.51) * v + np.float32(0.03))) / (v * (np.float32(2.43) * v + np.float32(0.59)) + np.float32(0.14))
    return np.power(np.clip(v, 0, 1), np.float32(1 / 2.4)).astype(np.float32)


def ref_extract(raw, channel, tonemap=False, max_exposure=1.0):
    """IImageEvaluator::ExtractColor -> (B,4,H,W) float32. channel: 0 mask, 1 normal, 2 depth, 3 colour."""
    raw = raw.astype(np.float32)
    out = np.empty((raw.shape[0], 4) + raw.shape[2:], np.float32)
    if channel == 3:
        out[:] = raw[:, :4]
        if tonemap:
            out[:, :3] = ref_tonemap(raw[:, :3], max_exposure)
    elif channel == 0:
        out[:, :3] = raw[:, 3:4]; out[:, 3] = 1.0
    elif channel == 1:
        out[:, :3] = raw[:, 4:7] * 0.5 + 0.5; out[:, 3] = raw[:, 3]
    else:
        mn, mx = np.min(raw[:, 7]), np.max(raw[:, 7])          # NaN propagates, as torch min / max
        out[:, :3] = (raw[:, 7] * (1 / (mx - mn)) + (-mn / (mx - mn)))[:, None]; out[:, 3] = 1.0
    return out


def ref_rgba8(rgba):
    """rgbaToInt per pixel -> (B,H,W) uint32."""
    q = np.clip(rgba.astype(np.float32) * np.float32(255), 0, 255).astype(np.uint32)      # unsigned() truncates
    return (q[:, 3] << 24) | (q[:, 2] << 16) | (q[:, 1] << 8) | q[:, 0]


@pytest.fixture()
def scene(env, plume_tf):
    g = Golden("c3_relu_gauss")
    vol = _volume(env, g.volnet_bytes)
    ie = _evaluator(env, vol, g.camera, plume_tf, 1 / 64)
    return g, ie


def test_extract_color_pass_matches_reference_kernels(env, scene):
    g, ie = scene
    raw_t = ie.render(56, 40)
    raw = raw_t.cpu().numpy()
    # depth is NaN on empty pixels (0/0 as in the reference): a frame without empty pixels for the depth channel
    filled = raw.copy(); filled[:, 7] = np.nan_to_num(filled[:, 7], nan=1.25)
    filled[:, 4:7] = np.random.default_rng(0).uniform(-1, 1, filled[:, 4:7].shape).astype(np.float32)   # the kernel is a pure function of the tensor
    filled_t = env.torch.from_numpy(filled).cuda()
    for ch, enum in ((0, env.pr.IImageEvaluator.Mask), (1, env.pr.IImageEvaluator.Normal), (2, env.pr.IImageEvaluator.Depth), (3, env.pr.IImageEvaluator.Color)):
        ie.selected_channel = enum
        got = ie.extract_color(filled_t).cpu().numpy()
        want = ref_extract(filled, ch)
        np.testing.assert_allclose(got, want, rtol=0, atol=2e-6, err_msg=str(enum))
        got8 = ie.extract_color(filled_t, output_format="rgba8").cpu().numpy().view(np.uint32)
        diff = np.abs((got8[..., None] >> np.arange(0, 32, 8) & 0xff).astype(int) - (ref_rgba8(want)[..., None] >> np.arange(0, 32, 8) & 0xff).astype(int))
        assert diff.max() <= (1 if ch in (1, 2) else 0), enum      # affine channels: one ulp of the fp32 value may cross an integer
        got16 = ie.extract_color(filled_t, output_format="rgba_f16").cpu().numpy()
        np.testing.assert_array_equal(got16, np.moveaxis(got, 1, -1).astype(np.float16))
    # ChannelDepth on the raw frame: NaN depth of empty pixels propagates through min / max like torch's (reference TODO, :86)
    ie.selected_channel = env.pr.IImageEvaluator.Depth
    assert np.isnan(ie.extract_color(raw_t).cpu().numpy()[:, :3]).all() == np.isnan(raw[:, 7]).any()
    # tonemapping (fixed and frame-derived exposure)
    ie.selected_channel = env.pr.IImageEvaluator.Color
    ie.use_tonemapping = True
    assert abs(ie.last_max_exposure - raw[:, :3].max()) < 1e-7
    for fixed in (False, True):
        ie.fix_max_exposure = fixed; ie.fixed_max_exposure = 0.37; ie.tonemapping_shoulder = 0.8
        e = max(0.001, (0.37 if fixed else raw[:, :3].max()) * 0.8)
        got = ie.extract_color(raw_t).cpu().numpy()
        want = ref_extract(raw, 3, True, e)
        np.testing.assert_allclose(got, want, rtol=0, atol=3e-4)        # div.approx / lg2.approx / ex2.approx of the fast-math build
        assert np.array_equal(got[:, 3], raw[:, 3])


@pytest.mark.parametrize("fmt", ["rgba_f32", "rgba8", "rgba_f16"])
def test_fused_colour_output_equals_render_then_extract(env, scene, fmt):
    g, ie = scene
    W, H = 56, 40
    raw_t = ie.render(W, H)
    for ch in (env.pr.IImageEvaluator.Color, env.pr.IImageEvaluator.Mask, env.pr.IImageEvaluator.Normal):
        ie.selected_channel = ch
        two_pass = ie.extract_color(raw_t, output_format=fmt).cpu().numpy()
        fused = ie.render_color(W, H, output_format=fmt).cpu().numpy()
        assert np.array_equal(fused, two_pass), (fmt, ch)
    ie.selected_channel = env.pr.IImageEvaluator.Color
    ie.use_tonemapping = True; ie.fix_max_exposure = True; ie.fixed_max_exposure = 0.5
    assert np.array_equal(ie.render_color(W, H, output_format=fmt).cpu().numpy(), ie.extract_color(raw_t, output_format=fmt).cpu().numpy())
    ie.fix_max_exposure = False
    with pytest.raises(RuntimeError, match="fix_max_exposure"):
        ie.render_color(W, H, output_format=fmt)
    ie.use_tonemapping = False
    ie.selected_channel = env.pr.IImageEvaluator.Depth
    with pytest.raises(RuntimeError, match="ChannelDepth"):
        ie.render_color(W, H, output_format=fmt)


@pytest.mark.parametrize("fmt,planes,dtype", [("raw8_f32", 8, np.float32), ("rgba_f32", 4, np.float32), ("rgba8", 1, np.uint32), ("rgba_f16", 1, np.float16)])
def test_host_entry_point_formats_and_band_streaming(env, scene, fmt, planes, dtype):
    """fvsrn_render_dvr_host: every output format, at a height that takes the four-band streamed path (H >= 256) and one that
    does not, into pageable and into pinned host memory; the host image equals the device render bit for bit."""
    g, ie = scene
    cam = np.ascontiguousarray(g.camera, np.float32); tf = np.ascontiguousarray(ie.ray_evaluator.tf._tensor().cpu().numpy())
    net = ie.volume.current_network(); net._ensure_prepared()
    for W, H in ((72, 264), (48, 40)):
        dev = (ie.render(W, H) if fmt == "raw8_f32" else ie.render_color(W, H, output_format=fmt)).cpu().numpy()
        strides = (planes * W * H, W * H, W, 1) if planes > 1 else (W * H, 0, W, 1)
        p = ie._params(W, H, env.torch.from_numpy(cam), env.torch.from_numpy(tf), strides)
        p.camera_matrix_dev = cam.ctypes.data; p.tf_dev = tf.ctypes.data
        p.output_format = env.pr.OUTPUT_FORMATS[fmt]; p.channel_mode = 3
        shape = (1, planes, H, W) if planes > 1 else ((1, H, W, 4) if fmt == "rgba_f16" else (1, H, W))
        out = np.zeros(shape, dtype)        # pageable host memory: staged through the device image, copied band by band
        env.lib.check(env.lib.lib().fvsrn_render_dvr_host(net._h, C.byref(p), out.ctypes.data, None))
        assert np.array_equal(np.nan_to_num(out.view(dev.dtype) if fmt == "rgba8" else out), np.nan_to_num(dev)), (fmt, W, H)
        # pinned (device-mapped) host memory: the packed formats are written by the kernel straight into the caller's buffer
        tdt = {np.float32: env.torch.float32, np.uint32: env.torch.int32, np.float16: env.torch.float16}[dtype]
        pinned = env.torch.zeros(shape, dtype=tdt).pin_memory()
        env.lib.check(env.lib.lib().fvsrn_render_dvr_host(net._h, C.byref(p), pinned.data_ptr(), None))
        assert np.array_equal(np.nan_to_num(pinned.numpy().view(dev.dtype)), np.nan_to_num(dev)), (fmt, W, H, "pinned")
