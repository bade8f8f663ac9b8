"""GPU parity tests of the fused DVR frame (ImageEvaluatorSimple.render) — see test_gpu_parity.py for the tolerances."""
import ctypes as C

import numpy as np
import pytest

from conftest import Golden, golden_names, psnr
from test_gpu_parity import env, _volume, _evaluator, RGBA_TOL  # noqa: F401

pytestmark = pytest.mark.gpu

DVR_CASES = [n for n in golden_names() if Golden(n).has_dvr]


def _check_steps(got_steps, want_steps, max_frac=0.02):
    """Step counts must equal the oracle's; a ray may differ only when its termination test (alpha vs 1-1e-5, or the
    density-min skip) is decided inside fp16 noise. Such rays are rare; they are counted and bounded."""
    diff = got_steps != want_steps
    assert diff.mean() <= max_frac, f"{diff.mean():.4f} of rays differ in step count"
    return float(diff.mean())


@pytest.mark.parametrize("name", DVR_CASES)
def test_dvr_matches_oracle_and_reference(env, name, plume_tf):
    g = Golden(name)
    vol = _volume(env, g.volnet_bytes)
    vol.current_network().set_time_and_ensemble(g.time, 0)
    ie = _evaluator(env, vol, g.camera, plume_tf, g.dvr_step)
    W = g.dvr_size
    img = ie.render(W, W)
    assert tuple(img.shape) == (1, 8, W, W) and img.dtype == env.torch.float32
    img = img.cpu().numpy()
    ref, steps, nvalid = env.oracle.render_dvr(g.net, width=W, height=W, camera=g.camera, fov_y=env.scenes.DEFAULT_FOV_Y,
                                               stepsize=g.dvr_step, tf=plume_tf, time=g.time)
    assert np.abs(img[:, :4] - ref[:, :4]).max() <= RGBA_TOL
    assert psnr(img[0, :4], g.dvr) >= 50.0                              # vs the reference's PyTorch marcher
    assert not img[:, 4:7].any()                                        # normals are zero for density networks
    # empty pixels: depth 0/0 = NaN, as the reference. A pixel whose absorption is exactly zero in the oracle can carry 1e-6 in the
    # kernel (fp16 accumulation moves a density across the first TF knot) and vice versa: such pixels are inside RGBA_TOL (asserted
    # above) and have no meaningful depth
    empty = (ref[0, 3] == 0) & (img[0, 3] == 0)
    assert np.isnan(img[0, 7][empty]).all() and (ref[0, 3] == 0).sum() - empty.sum() <= 0.001 * empty.size
    both = (ref[0, 3] > 1e-4) & (img[0, 3] > 1e-4)
    np.testing.assert_allclose(img[0, 7][both], ref[0, 7][both], rtol=0, atol=5e-3)
    _check_steps(ie.last_step_counts.cpu().numpy(), steps)
    counters = ie.last_counters.cpu().numpy()
    assert abs(int(counters[0]) - nvalid) <= 0.001 * nvalid and counters[1] >= counters[0]
    color = ie.extract_color(env.torch.from_numpy(img).cuda()).cpu().numpy()
    assert np.array_equal(np.nan_to_num(color), np.nan_to_num(img[:, :4]))


def test_dvr_no_early_out_counts_are_exact(env, plume_tf):
    """Without early termination the step count depends only on fp32 ray setup -> bit-exact vs the oracle."""
    g = Golden("c3_relu_gauss")
    vol = _volume(env, g.volnet_bytes)
    ie = _evaluator(env, vol, g.camera, plume_tf, 1 / 100, early_out=False)
    ie.render(40, 24)
    _, steps, _ = env.oracle.render_dvr(g.net, width=40, height=24, camera=g.camera, fov_y=env.scenes.DEFAULT_FOV_Y,
                                        stepsize=1 / 100, tf=plume_tf, early_out=False)
    assert np.array_equal(ie.last_step_counts.cpu().numpy(), steps)


def test_dvr_texture_tf_alpha_blend_density_range(env, plume_tf):
    g = Golden("sine_48")
    tex = env.scenes.texture_tf_from_piecewise(plume_tf, 256)
    vol = _volume(env, g.volnet_bytes)
    ie = _evaluator(env, vol, g.camera, tex, 1 / 80, tf_kind=1, min_density=0.2, max_density=0.9)
    ie.ray_evaluator.blending.blendMode = env.pr.Blending.Alpha
    img = ie.render(32, 32).cpu().numpy()
    ref, steps, _ = env.oracle.render_dvr(g.net, width=32, height=32, camera=g.camera, fov_y=env.scenes.DEFAULT_FOV_Y, stepsize=1 / 80,
                                          tf=tex, tf_kind=1, density_min=0.2, density_max=0.9, blend_mode=0)
    # density_min = 0.2 makes the image a DISCONTINUOUS function of the network output: a sample whose density lies within
    # rounding noise of the threshold is either skipped or composited, which moves its pixel by up to that sample's whole
    # contribution (2e-3 here). No tolerance on the network output bounds that; the gate is therefore 2e-3 on every pixel
    # but at most two of the 1024 (threshold flips, bounded by 3e-3). The reference-exact mode has no such pixels: it is
    # compared bit for bit with the reference kernel (test_gpu_reference_exact.py).
    d = np.abs(img[:, :4] - ref[:, :4]).max(axis=(0, 1))
    assert (d > RGBA_TOL).sum() <= 2 and d.max() <= 1.5 * RGBA_TOL, (int((d > RGBA_TOL).sum()), float(d.max()))
    _check_steps(ie.last_step_counts.cpu().numpy(), steps)


@pytest.mark.parametrize("points", [2, 3, 97, 204])
def test_dvr_long_irregular_piecewise_tf(env, points):
    """Piecewise TFs from the minimum of 2 control points to the maximum of 204, with clustered and duplicated positions:
    the interval the kernels pick (tcgen05: binned start index + scan; mma.sync: linear scan) must be the one the
    reference's linear search picks (renderer_tf_piecewise.cuh:27-52), for every density."""
    rng = np.random.default_rng(points)
    pos = np.sort(np.concatenate([rng.random(points // 2) * 0.1 + 0.3, rng.random(points - points // 2)])).astype(np.float32)
    if points >= 8:
        pos[5] = pos[4]                       # a zero-width interval
        pos[points // 2:points // 2 + 3] = pos[points // 2]
    pos[0], pos[-1] = 0.0, 1.0
    # colour / absorption are smooth functions of the position, sampled at the irregular control points (a TF with random
    # values per point, or non-zero absorption at density 0, would amplify the fp16 noise of the density itself and test
    # the network, not the search); absorption 0 at density 0 keeps the image continuous across the density-min skip
    tf = np.zeros((1, points, 5), np.float32)
    tf[0, :, 0] = 0.5 + 0.5 * np.sin(3.0 * pos); tf[0, :, 1] = pos; tf[0, :, 2] = 1.0 - pos * pos
    tf[0, :, 3] = 20.0 * pos * pos; tf[0, :, 4] = pos
    g = Golden("c3_relu_gauss")
    vol = _volume(env, g.volnet_bytes)
    ie = _evaluator(env, vol, g.camera, tf, 1 / 64)
    img = ie.render(48, 40).cpu().numpy()
    ref, steps, _ = env.oracle.render_dvr(g.net, width=48, height=40, camera=g.camera, fov_y=env.scenes.DEFAULT_FOV_Y, stepsize=1 / 64, tf=tf)
    assert np.abs(img[:, :4] - ref[:, :4]).max() <= RGBA_TOL
    _check_steps(ie.last_step_counts.cpu().numpy(), steps)


def test_dvr_rgbo_network_skips_tf(env):
    g = Golden("rgbo_32")
    cam = env.scenes.camera_reference_frame(**env.scenes.DEFAULT_CAMERA)
    vol = _volume(env, g.volnet_bytes)
    ie = _evaluator(env, vol, cam, None, 1 / 64, tf_kind=-1)
    img = ie.render(32, 32).cpu().numpy()
    ref, _, _ = env.oracle.render_dvr(g.net, width=32, height=32, camera=cam, fov_y=env.scenes.DEFAULT_FOV_Y, stepsize=1 / 64, tf=None)
    assert np.abs(img[:, :4] - ref[:, :4]).max() <= RGBA_TOL


@pytest.mark.parametrize("dir_in_fourier", [False, True])
def test_dvr_network_with_direction_input(env, plume_tf, dir_in_fourier):
    """The DIR kernels in the DVR frame: the ray direction is a network input (channels 4..6), optionally also inside the
    Fourier features (USE_DIRECTION 1 / 2, renderer_volume_tensorcores.cuh:768-807) - per ray constant, per pixel different."""
    net = env.scenes.make_network(hidden=32, num_layers=3, grid_res=8, has_direction=True, fourier="gauss", fourier_std=0.7,
                                  direction_in_fourier=dir_in_fourier, seed=5, last_gain=4.0)
    assert net.direction_in_fourier == dir_in_fourier
    cam = env.scenes.camera_reference_frame(**env.scenes.DEFAULT_CAMERA)
    vol = _volume(env, net)
    tf = plume_tf.copy(); tf[..., 3] *= 0.25         # this random network sits high in the TF: keep the frame translucent like the BASELINE scenes
    ie = _evaluator(env, vol, cam, tf, 1 / 64)
    img = ie.render(48, 32).cpu().numpy()
    ref, steps, _ = env.oracle.render_dvr(net, width=48, height=32, camera=cam, fov_y=env.scenes.DEFAULT_FOV_Y, stepsize=1 / 64, tf=tf)
    assert 0.05 < ref[:, 3].max() < 0.95
    assert np.abs(img[:, :4] - ref[:, :4]).max() <= RGBA_TOL
    _check_steps(ie.last_step_counts.cpu().numpy(), steps)


def test_dvr_per_view_transfer_functions_with_tiles_and_times(env, plume_tf):
    """tf_batch == B (one TF per view, re-staged per tile in the tcgen05 kernel) together with per-view latent times and a
    screen tile - the combination a sharded time sweep with per-view TFs makes."""
    g = Golden("c5_time")
    cams = env.scenes.camera_sweep(3)
    tfs = np.concatenate([plume_tf, plume_tf, plume_tf], 0).copy()
    tfs[1, :, 3] *= 0.4; tfs[2, :, 3] *= 2.5; tfs[2, :, :3] = tfs[2, :, :3][:, ::-1]
    times = [0.25, 1.0, 1.75]
    W, H, tile = 64, 40, (8, 4, 40, 24)
    vol = _volume(env, g.volnet_bytes)
    ie = _evaluator(env, vol, cams, tfs, 1 / 64)
    ie.view_times = times
    ie.tile = tile
    img = ie.render(W, H).cpu().numpy()
    x0, y0, w, h = tile
    for b, t in enumerate(times):
        ref, _, _ = env.oracle.render_dvr(g.net, width=W, height=H, camera=cams[b:b + 1], fov_y=env.scenes.DEFAULT_FOV_Y,
                                          stepsize=1 / 64, tf=tfs[b:b + 1], time=t, tile=tile)
        assert np.abs(img[b, :4, y0:y0 + h, x0:x0 + w] - ref[0, :4, y0:y0 + h, x0:x0 + w]).max() <= RGBA_TOL, (b, t)
    assert not np.array_equal(img[0, :4], img[1, :4])
    # without per-view times: one launch over the batch, every view with its own TF
    ie.view_times = None
    vol.current_network().set_time_and_ensemble(1.0, 0)
    img = ie.render(W, H).cpu().numpy()
    for b in range(3):
        ref, _, _ = env.oracle.render_dvr(g.net, width=W, height=H, camera=cams[b:b + 1], fov_y=env.scenes.DEFAULT_FOV_Y,
                                          stepsize=1 / 64, tf=tfs[b:b + 1], time=1.0, tile=tile)
        assert np.abs(img[b, :4, y0:y0 + h, x0:x0 + w] - ref[0, :4, y0:y0 + h, x0:x0 + w]).max() <= RGBA_TOL, b


def test_dvr_tiles_and_views_compose(env, plume_tf):
    """Screen-tile sharding (C4) and view batches (C5): tiles rendered separately equal the full frame bit for bit."""
    g = Golden("c1_relu_nerf")
    cams = env.scenes.camera_sweep(3)
    vol = _volume(env, g.volnet_bytes)
    ie = _evaluator(env, vol, cams, plume_tf, 1 / 64)
    W, H = 64, 40
    full = ie.render(W, H).cpu().numpy()
    assert full.shape == (3, 8, H, W)
    acc = np.zeros_like(full)
    for (x0, y0, w, h) in [(0, 0, 24, 40), (24, 0, 40, 16), (24, 16, 40, 24)]:
        ie.tile = (x0, y0, w, h)
        part = ie.render(W, H).cpu().numpy()
        acc[:, :, y0:y0 + h, x0:x0 + w] = part[:, :, y0:y0 + h, x0:x0 + w]
    assert np.array_equal(np.nan_to_num(acc), np.nan_to_num(full))
    ie.tile = None
    for b in range(3):
        ie.camera.set_parameters(env.torch.from_numpy(cams[b:b + 1]))
        one = ie.render(W, H).cpu().numpy()
        assert np.array_equal(np.nan_to_num(one[0]), np.nan_to_num(full[b]))


@pytest.mark.parametrize("W,H", [(50, 30), (37, 19), (1, 1)])
def test_dvr_ragged_image_sizes(env, plume_tf, W, H):
    """Image sizes that are no multiple of the screen tile (16x8 / 8x4): partial tiles must neither write outside the
    image nor change the pixels inside it."""
    g = Golden("c3_relu_gauss")
    vol = _volume(env, g.volnet_bytes)
    ie = _evaluator(env, vol, g.camera, plume_tf, 1 / 64)
    guard = env.torch.full((1, 8, H + 2, W + 2), 7.0, dtype=env.torch.float32, device="cuda")
    view = guard[:, :, 1:H + 1, 1:W + 1]
    p = ie._params(W, H, ie.camera.get_parameters(), ie.ray_evaluator.tf._tensor(), view.stride())
    sn = vol.current_network(); sn._ensure_prepared()
    stream = C.c_void_p(env.torch.cuda.current_stream().cuda_stream)
    env.lib.check(env.lib.lib().fvsrn_render_dvr(sn._h, C.byref(p), view.data_ptr(), stream))
    env.torch.cuda.synchronize()
    got = guard.cpu().numpy()
    border = np.ones((H + 2, W + 2), bool); border[1:H + 1, 1:W + 1] = False
    assert (got[0][:, border] == 7.0).all()                               # nothing written outside the image
    ref, steps, _ = env.oracle.render_dvr(g.net, width=W, height=H, camera=g.camera, fov_y=env.scenes.DEFAULT_FOV_Y, stepsize=1 / 64,
                                          tf=plume_tf)
    assert np.abs(got[:, :4, 1:H + 1, 1:W + 1] - ref[:, :4]).max() <= RGBA_TOL


def test_dvr_ragged_tiles_equal_the_full_frame(env, plume_tf):
    """Rectangles that start and end off the 16x8 / 8x4 tile raster reproduce the full frame's pixels bit for bit."""
    g = Golden("c3_relu_gauss")
    vol = _volume(env, g.volnet_bytes)
    ie = _evaluator(env, vol, g.camera, plume_tf, 1 / 64)
    W, H = 53, 31
    full = ie.render(W, H).cpu().numpy()
    for (x0, y0, w, h) in [(3, 5, 17, 3), (0, 0, 1, 1), (52, 30, 1, 1), (7, 1, 46, 30), (16, 8, 16, 8)]:
        ie.tile = (x0, y0, w, h)
        part = ie.render(W, H).cpu().numpy()
        assert np.array_equal(np.nan_to_num(part[:, :, y0:y0 + h, x0:x0 + w]), np.nan_to_num(full[:, :, y0:y0 + h, x0:x0 + w])), (x0, y0, w, h)
        outside = np.ones((H, W), bool); outside[y0:y0 + h, x0:x0 + w] = False
        assert not np.nan_to_num(part[0][:, outside]).any()                   # a tile render leaves the rest of the (zeroed) frame alone
    ie.tile = None


def test_dvr_per_view_times(env, plume_tf):
    g = Golden("c5_time")
    cams = env.scenes.camera_sweep(4)
    times = [0.0, 0.5, 1.25, 2.0]
    vol = _volume(env, g.volnet_bytes)
    ie = _evaluator(env, vol, cams, plume_tf, 1 / 64)
    ie.view_times = times
    img = ie.render(32, 32).cpu().numpy()
    for b, t in enumerate(times):
        ref, _, _ = env.oracle.render_dvr(g.net, width=32, height=32, camera=cams[b:b + 1], fov_y=env.scenes.DEFAULT_FOV_Y,
                                          stepsize=1 / 64, tf=plume_tf, time=t)
        assert np.abs(img[b, :4] - ref[0, :4]).max() <= RGBA_TOL, (b, t)


@pytest.mark.parametrize("encoding", [1, 2])
def test_dvr_byte_encoded_time_keyframes(env, plume_tf, encoding):
    """Byte-encoded latent grids stay 8-bit on the device and are decoded after the trilinear filter; with two keyframes
    each is decoded with its OWN offset / scale (the reference kernel uses A's for both; oracle b_uses_a=False restates the
    fixed behaviour, the reference-exact mode the reference's)."""
    net = env.scenes.encode_grids(env.scenes.make_network(hidden=32, num_layers=4, grid_res=16, seed=5, last_gain=6.0, time_keyframes=3), encoding)
    assert net.time_grids[0].encoding == encoding and net.time_grids[0].raw is not None
    cam = env.scenes.camera_reference_frame(**env.scenes.DEFAULT_CAMERA)
    vol = _volume(env, net)
    vol.current_network().set_time_and_ensemble(1.4, 0)
    ie = _evaluator(env, vol, cam, plume_tf, 1 / 64)
    img = ie.render(40, 32).cpu().numpy()
    ref, steps, _ = env.oracle.render_dvr(net, width=40, height=32, camera=cam, fov_y=env.scenes.DEFAULT_FOV_Y, stepsize=1 / 64, tf=plume_tf, time=1.4)
    assert np.abs(img[:, :4] - ref[:, :4]).max() <= RGBA_TOL
    _check_steps(ie.last_step_counts.cpu().numpy(), steps)
    pos = np.random.default_rng(3).random((2000, 3), dtype=np.float32)
    got = vol.evaluate(env.torch.from_numpy(pos).cuda()).cpu().numpy()
    assert np.abs(got - env.oracle.evaluate(net, pos, time=1.4)).max() < 4e-3


def test_dvr_supersampling_matches_oracle(env, plume_tf):
    g = Golden("c3_relu_gauss")
    vol = _volume(env, g.volnet_bytes)
    ie = _evaluator(env, vol, g.camera, plume_tf, 1 / 64)
    ie.samples_per_pixel = 4
    img = ie.render(24, 24).cpu().numpy()
    ref, _, _ = env.oracle.render_dvr(g.net, width=24, height=24, camera=g.camera, fov_y=env.scenes.DEFAULT_FOV_Y, stepsize=1 / 64,
                                      tf=plume_tf, spp=4)
    assert np.abs(img[:, :4] - ref[:, :4]).max() <= RGBA_TOL
    ie.samples_per_pixel = 3
    with pytest.raises(RuntimeError, match="square"):
        ie.render(24, 24)


def test_dvr_host_entry_point_and_strides(env, plume_tf):
    g = Golden("c1_relu_nerf")
    vol = _volume(env, g.volnet_bytes)
    ie = _evaluator(env, vol, g.camera, plume_tf, 1 / 64)
    W = H = 48
    dev = ie.render(W, H).cpu().numpy()
    net = vol.current_network()
    cam = np.ascontiguousarray(g.camera, np.float32); tf = np.ascontiguousarray(plume_tf, np.float32)
    p = ie._params(W, H, env.torch.from_numpy(cam), env.torch.from_numpy(tf), (8 * W * H, W * H, W, 1))
    p.camera_matrix_dev = cam.ctypes.data; p.tf_dev = tf.ctypes.data
    out = np.empty((1, 8, H, W), np.float32)
    env.lib.check(env.lib.lib().fvsrn_render_dvr_host(net._h, C.byref(p), out.ctypes.data, None))
    assert np.array_equal(np.nan_to_num(out), np.nan_to_num(dev))
    # channel-last output tensor: any stride layout is accepted (renderer/pytorch_utils.h:39-43)
    buf = env.torch.empty((1, H, W, 8), device="cuda").permute(0, 3, 1, 2)
    got = ie.render(W, H, out=buf)
    assert got.data_ptr() == buf.data_ptr()
    assert np.array_equal(np.nan_to_num(got.cpu().numpy()), np.nan_to_num(dev))


def test_full_size_properties(env, plume_tf):
    """BASELINE C3 size (1024^2, step 1/512): properties that need no oracle run — determinism, tile composition,
    alpha in [0,1], empty pixels exactly where the ray misses the box, useful <= executed samples."""
    net = env.scenes.make_network(hidden=32, num_layers=4, grid_res=32, activation="ReLU", seed=42, last_gain=6.0)
    cam = env.scenes.camera_reference_frame(**env.scenes.DEFAULT_CAMERA)
    vol = _volume(env, net)
    ie = _evaluator(env, vol, cam, plume_tf, 1 / 512)
    a = ie.render(1024, 1024).cpu().numpy()
    c1 = ie.last_counters.cpu().numpy().copy()
    b = ie.render(1024, 1024).cpu().numpy()
    assert np.array_equal(np.nan_to_num(a), np.nan_to_num(b))
    assert a[:, 3].min() >= 0 and a[:, 3].max() <= 1.0 and np.isfinite(a[:, :4]).all()
    steps = ie.last_step_counts.cpu().numpy()
    assert (a[:, 3][steps == 0] == 0).all() and steps.max() <= 887 and (steps > 0).mean() > 0.2
    assert 0 < c1[0] <= c1[1] and c1[0] == steps.sum()
    ie.tile = (512, 256, 256, 512)
    t = ie.render(1024, 1024).cpu().numpy()
    assert np.array_equal(np.nan_to_num(t[:, :, 256:768, 512:768]), np.nan_to_num(a[:, :, 256:768, 512:768]))
