"""
Ground-truth density grid through the same marcher: fvsrn_render_dvr_grid / fvsrn_evaluate_grid against the
REFERENCE's own DVR kernel compiled with VolumeInterpolationGrid (oracle/build_ref_ptx.py grid_source, run through
oracle/ref_kernel.py ReferenceGridDvrKernel) - bit for bit: RGBA, depth, per-ray step counts - and against a NumPy
trilinear restatement of renderer_volume_grid.cuh:110-139, 196-231.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def blob_volume(n=48, dtype=np.float32, seed=0):
    """smooth synthetic density in [0,1]: a few gaussian blobs (Z,Y,X)"""
    rng = np.random.default_rng(seed)
    z, y, x = np.meshgrid(*(np.linspace(0, 1, n),) * 3, indexing="ij")
    d = np.zeros((n, n, n))
    for _ in range(6):
        c = rng.uniform(0.2, 0.8, 3); s = rng.uniform(0.05, 0.2)
        d += rng.uniform(0.3, 1.0) * np.exp(-((x - c[0]) ** 2 + (y - c[1]) ** 2 + (z - c[2]) ** 2) / (2 * s * s))
    d = np.clip(d / d.max(), 0, 1)
    return (d * 255).astype(np.uint8) if dtype == np.uint8 else d.astype(np.float32)


def make(pr, torch, scenes, vol, tf, step, world=(1.0, 1.0, 1.0)):
    if vol.dtype == np.uint8:       # a uchar feature goes through the .cvol container ("CVOL" v1, uncompressed)
        import struct
        Z, Y, X = vol.shape
        data = b"CVOL" + struct.pack("<i3fiii", 1, *world, 1, 0, 0) + struct.pack("<i", 7) + b"density" + struct.pack("<QQQii", X, Y, Z, 1, 0) + vol.tobytes()
        v = pr.Volume.from_bytes(data)
    else:
        v = pr.Volume.from_numpy(vol, world)
    grid = pr.VolumeInterpolationGrid(); grid.setSource(v)
    ie = pr.ImageEvaluatorSimple(); ie.volume = grid
    ie.camera.set_parameters(torch.from_numpy(scenes.camera_reference_frame(**scenes.DEFAULT_CAMERA)))
    ie.ray_evaluator.tf.tensor.value = torch.from_numpy(tf).cuda()
    ie.ray_evaluator.stepsize = step
    ie.collect_statistics = True
    return ie, grid


@pytest.mark.parametrize("dtype,world,absorb", [(np.float32, (1.0, 1.0, 1.0), 1.0), (np.uint8, (1.0, 1.0, 1.0), 8.0), (np.float32, (1.0, 0.5, 0.75), 8.0)])
def test_grid_dvr_is_bit_identical_to_the_reference_kernel(plume_tf, dtype, world, absorb):
    import torch
    import ref_kernel
    import fvsrn_b200  # noqa: F401
    from fvsrn_b200 import pyrenderer as pr, scenes
    if not ref_kernel.available():
        pytest.skip("oracle/_ref/*.ptx not built")
    vol = blob_volume(48, dtype)
    tf = plume_tf.copy(); tf[..., 3] *= absorb
    W, H, step = 256, 192, 1 / 256
    ie, _ = make(pr, torch, scenes, vol, tf, step, world)
    img = ie.render(W, H).cpu().numpy(); steps = ie.last_step_counts.cpu().numpy()
    cam = scenes.camera_reference_frame(**scenes.DEFAULT_CAMERA)
    want = ref_kernel.ReferenceGridDvrKernel(vol, world).render(W, H, cam, scenes.DEFAULT_FOV_Y, step, tf).cpu().numpy()
    counted = ref_kernel.ReferenceGridDvrKernel(vol, world, patched_steps=True).render(W, H, cam, scenes.DEFAULT_FOV_Y, step, tf).cpu().numpy()
    assert np.array_equal(want[:, :4], counted[:, :4])
    assert want[:, 3].max() > 0.5 and (absorb == 1.0 or (want[:, 3] >= 1 - 1e-5).any())     # content; saturating rays when opaque
    a, b = np.nan_to_num(img, nan=-7.0), np.nan_to_num(want, nan=-7.0)
    assert np.array_equal(steps[0], np.rint(counted[0, 4]).astype(np.int32))
    assert np.array_equal(a[:, :4], b[:, :4]) and np.array_equal(a[:, 7], b[:, 7]) and not img[:, 4:7].any()
    # fused colour output of the grid renderer equals render + extract_color
    fused = ie.render_color(W, H, output_format="rgba8").cpu().numpy()
    assert np.array_equal(fused, ie.extract_color(torch.from_numpy(img).cuda(), output_format="rgba8").cpu().numpy())


def test_grid_evaluate_matches_trilinear_restatement():
    import torch
    import fvsrn_b200  # noqa: F401
    from fvsrn_b200 import pyrenderer as pr
    vol = blob_volume(20, np.float32, seed=3)
    grid = pr.VolumeInterpolationGrid(); grid.setSource(pr.Volume.from_numpy(vol))
    rng = np.random.default_rng(1)
    pos = np.concatenate([rng.random((4000, 3)), rng.uniform(-0.2, 1.2, (500, 3))]).astype(np.float32)
    got = grid.evaluate(torch.from_numpy(pos).cuda()).cpu().numpy()[:, 0]
    # texture fetch at unnormalised coordinate p*res: texel centres at i + 0.5, clamp addressing, linear filter
    n = vol.shape[0]
    u = pos.astype(np.float64) * n - 0.5
    i0 = np.floor(u).astype(int); f = u - i0
    want = np.zeros(len(pos))
    for dz in (0, 1):
        for dy in (0, 1):
            for dx in (0, 1):
                xi = np.clip(i0[:, 0] + dx, 0, n - 1); yi = np.clip(i0[:, 1] + dy, 0, n - 1); zi = np.clip(i0[:, 2] + dz, 0, n - 1)
                w = (f[:, 0] if dx else 1 - f[:, 0]) * (f[:, 1] if dy else 1 - f[:, 1]) * (f[:, 2] if dz else 1 - f[:, 2])
                want += w * vol[zi, yi, xi]
    assert np.abs(got - want).max() < 4e-3          # 8-bit fractional weights of the texture unit
    assert tuple(grid.box_min()) == (-0.5, -0.5, -0.5)
    with pytest.raises(RuntimeError):
        pr.VolumeInterpolationGrid().evaluate(torch.zeros((4, 3), device="cuda"))
