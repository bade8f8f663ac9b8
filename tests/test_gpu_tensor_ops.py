"""
The tensor operators beside the fused path: camera.generate_rays and tf.evaluate[_with_previous] - what the
reference's PyTorch marcher (applications/volnet/raytracing.py:103, 150-154) calls - against NumPy restatements of
renderer_camera_kernels.cuh:12-45 + renderer_camera.cuh:34-52 and renderer_tf_kernels.cuh:11-70 + renderer_tf_piecewise.cuh:
27-62 / renderer_tf_texture.cuh:46-56 (the same stand-ins tests/golden/make_golden.py feeds to the reference's raytracing.py).
Tolerance: the reference kernels are fast-math builds (div.approx, sin/cos.approx, rsqrt.approx): a few fp32 ulp.
"""
import numpy as np
import pytest

from test_gpu_parity import env  # noqa: F401

pytestmark = pytest.mark.gpu


def ref_rays(cam, fov_y, W, H):
    cam = cam.astype(np.float64)
    out_s = np.empty((cam.shape[0], H, W, 3)); out_d = np.empty_like(out_s)
    for b, (eye, right, up) in enumerate(cam):
        front = np.cross(up, right)
        ty = np.tan(fov_y / 2); tx = ty * (W / H)
        nx = 2 * (np.arange(W) + 0.5) / W - 1; ny = 2 * (np.arange(H) + 0.5) / H - 1
        d = front[None, None] + (nx * tx)[None, :, None] * right[None, None] + (ny * ty)[:, None, None] * up[None, None]
        out_d[b] = d / np.linalg.norm(d, axis=-1, keepdims=True); out_s[b] = eye
    return out_s, out_d


def ref_tf(kind, tf, d, dmin, dmax, step):
    tf = tf[0].astype(np.float64); d = d.astype(np.float64)
    out = np.zeros((d.shape[0], 4))
    on = d >= dmin
    d2 = np.clip((d - dmin) / (dmax - dmin), 0, 1)
    R = tf.shape[0]
    if kind == 0:
        pos = tf[:, 4]
        idx = np.array([next((i for i in range(R - 2) if pos[i + 1] > v), R - 2) for v in d2])
        p0, p1 = pos[idx], pos[idx + 1]
        f = (np.clip(d2, p0, p1) - p0) / (p1 - p0)
        out = tf[idx, :4] + f[:, None] * (tf[idx + 1, :4] - tf[idx, :4])
    else:
        x = d2 * R - 0.5; fl = np.floor(x); df = x - fl
        i0 = np.clip(fl.astype(int), 0, R - 1); i1 = np.clip(fl.astype(int) + 1, 0, R - 1)
        out = tf[i0] + df[:, None] * (tf[i1] - tf[i0])
    out[:, 3] *= step
    out[~on] = 0
    return out


def test_generate_rays(env):
    cams = env.scenes.camera_sweep(3)
    cam = env.pr.CameraOnASphere(); cam.set_parameters(env.torch.from_numpy(cams))
    for W, H in ((64, 48), (33, 70)):
        s, d = cam.generate_rays(W, H)
        ws, wd = ref_rays(cams, cam.fovY_radians, W, H)
        assert tuple(s.shape) == (3, H, W, 3) and tuple(d.shape) == (3, H, W, 3)
        np.testing.assert_allclose(s.cpu().numpy(), ws, rtol=0, atol=1e-7)
        np.testing.assert_allclose(d.cpu().numpy(), wd, rtol=0, atol=2e-6)
    with pytest.raises(RuntimeError):
        cam.generate_rays(8, 8, True)


def test_evaluate_tf(env, plume_tf):
    torch = env.torch
    rng = np.random.default_rng(0)
    d = np.concatenate([rng.uniform(-0.2, 1.3, 5000), [0.0, 1.0, 0.1, 0.9, plume_tf[0, 2, 4]]]).astype(np.float32)
    dt = torch.from_numpy(d[:, None]).cuda()
    tfp = env.pr.TransferFunctionPiecewiseLinear(); tfp.tensor.value = torch.from_numpy(plume_tf).cuda()
    tft = env.pr.TransferFunctionTexture(); tft.tensor.value = torch.from_numpy(env.scenes.texture_tf_from_piecewise(plume_tf, 256)).cuda()
    for kind, tf, arr in ((0, tfp, plume_tf), (1, tft, tft.tensor.value.cpu().numpy())):
        for dmin, dmax in ((0.0, 1.0), (0.1, 0.9)):
            got = tf.evaluate(dt, dmin, dmax).cpu().numpy()
            want = ref_tf(kind, arr, d, dmin, dmax, 1.0)
            # interval boundaries: a density within an ulp of a control point may pick the neighbouring interval - same value (continuous TF)
            np.testing.assert_allclose(got, want, rtol=2e-5, atol=2e-5)
            got = tf.evaluate_with_previous(dt, dmin, dmax, dt, 1 / 512).cpu().numpy()
            np.testing.assert_allclose(got, ref_tf(kind, arr, d, dmin, dmax, 1 / 512), rtol=2e-5, atol=2e-6)
    with pytest.raises(RuntimeError):
        tfp.evaluate(torch.zeros((4, 1)), 0.0, 1.0)
