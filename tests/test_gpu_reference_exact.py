"""
The reference-exact kernel mode (csrc/kernels_exact.cuh) against the reference's own tensor-core kernel on the same B200:
north_star's third criterion - "early-termination step counts bit-exact against the reference's tensor-core kernel when
it is run with the same precision". Gated bit for bit: RGBA, depth and per-ray step counts, early termination ON, at the
BASELINE configs' own sizes (C3 1024^2 step 1/512, ReLU / SnakeAlt, NeRF / Gaussian Fourier, piecewise / texture TF,
translucent and saturating frames; C4 64x6 network; C5 time-interpolated keyframes) and on the committed golden scenes.
"""
import numpy as np
import pytest

from conftest import Golden
from test_gpu_baseline_configs import CASES, ref, make_case_network  # noqa: F401

pytestmark = pytest.mark.gpu

EXACT_CASES = [n for n in CASES]
GOLDEN_CASES = ["c3_relu_gauss", "c3_snakealt_nerf", "c1_relu_nerf"]


def _assert_identical(img, steps, want, want_steps, what):
    a, b = np.nan_to_num(img, nan=-7.0), np.nan_to_num(want, nan=-7.0)       # depth is NaN on empty pixels in both
    assert np.array_equal(steps, want_steps), (what, int((steps != want_steps).sum()), "rays differ in step count")
    assert np.array_equal(a[:, :4], b[:, :4]), (what, int((a[:, :4] != b[:, :4]).any(1).sum()), "pixels differ in RGBA")
    assert np.array_equal(a[:, 7], b[:, 7]), (what, "depth differs")
    assert not img[:, 4:7].any()


@pytest.mark.parametrize("name", EXACT_CASES)
def test_reference_exact_mode_is_bit_identical_at_baseline_configs(ref, name, plume_tf):
    import parity
    import fvsrn_b200  # noqa: F401
    from fvsrn_b200 import scenes, volnet
    kw, size, step, tf_kind, time = CASES[name][:5]
    absorption = CASES[name][5] if len(CASES[name]) > 5 else 1.0
    net = make_case_network(kw)
    cam = scenes.camera_reference_frame(**scenes.DEFAULT_CAMERA)
    tfp = plume_tf.copy(); tfp[..., 3] *= absorption
    tf = scenes.texture_tf_from_piecewise(tfp, 256) if tf_kind == 1 else tfp
    want, want_steps, _ = parity.render_reference(net, size, size, cam, scenes.DEFAULT_FOV_Y, tf, step, tf_kind=tf_kind,
                                                  time=0.0 if time is None else time)
    img, steps = parity.render_ours(volnet.dumps(net), "reference_exact", size, size, cam, tf, step, tf_kind=tf_kind, time=time)
    _assert_identical(img, steps, want, want_steps, name)


@pytest.mark.parametrize("name", GOLDEN_CASES)
def test_reference_exact_mode_is_bit_identical_on_golden_scenes(ref, name, plume_tf):
    import parity
    from fvsrn_b200 import scenes
    g = Golden(name)
    W = g.dvr_size
    want, want_steps, _ = parity.render_reference(g.net, W, W, g.camera, scenes.DEFAULT_FOV_Y, plume_tf, g.dvr_step)
    img, steps = parity.render_ours(g.volnet_bytes, "reference_exact", W, W, g.camera, plume_tf, g.dvr_step)
    _assert_identical(img, steps, want, want_steps, name)


def test_ray_setup_is_bit_identical_in_every_kernel_mode(ref, plume_tf):
    """Without early termination the step count depends only on the fp32 ray setup, which every kernel family shares with
    the reference-exact mode (kernels.cuh make_ray): counts equal the reference kernel's for every ray of a 1024^2 frame."""
    import parity
    from fvsrn_b200 import scenes, volnet
    kw, size, step = CASES["C3_relu_nerf"][:3]
    net = make_case_network(kw)
    cam = scenes.camera_reference_frame(**scenes.DEFAULT_CAMERA)
    want, want_steps, _ = parity.render_reference(net, size, size, cam, scenes.DEFAULT_FOV_Y, plume_tf, step)
    assert want[:, 3].max() < 0.9                           # translucent frame: no ray terminates early
    for mode in ("tcgen05", "mma_sync", "reference_exact"):
        _, steps = parity.render_ours(volnet.dumps(net), mode, size, size, cam, plume_tf, step)
        assert np.array_equal(steps, want_steps), mode


def test_reference_exact_mode_rejects_what_it_does_not_cover(plume_tf):
    import torch
    from fvsrn_b200 import pyrenderer as pr, scenes, volnet
    net = scenes.make_network(hidden=32, num_layers=4, grid_res=8, output_mode="rgbo", seed=1)
    sn = pr.SceneNetwork.from_bytes(volnet.dumps(net)); sn.kernel_mode = "reference_exact"
    vol = pr.VolumeInterpolationNetwork(); vol.set_network(sn)
    ie = pr.ImageEvaluatorSimple(); ie.volume = vol
    ie.camera.set_parameters(torch.from_numpy(scenes.camera_reference_frame(**scenes.DEFAULT_CAMERA)))
    with pytest.raises(RuntimeError, match="reference-exact"):
        ie.render(32, 32)
    with pytest.raises(RuntimeError, match="reference-exact"):
        vol.evaluate(torch.rand((64, 3), device="cuda"))


def test_alpha_blending_without_early_out_and_a_density_window(ref, plume_tf):
    """The other values of the marcher's compile-time switches - BLENDING_MODE alpha (renderer_blending.cuh:40-44),
    RAY_EVALUATION_STEPPING__ENABLE_EARLY_OUT 0 - and a density window [0.05, 0.6] (samples below it are skipped, above it
    clamped): reference-exact mode bit-identical, fast modes within the image gate."""
    import parity
    from fvsrn_b200 import scenes, volnet
    kw, size, step = CASES["C3_relu_nerf"][:3]
    size = 512
    net = make_case_network(kw)
    cam = scenes.camera_reference_frame(**scenes.DEFAULT_CAMERA)
    rng = (0.05, 0.6)
    want, want_steps, _ = parity.render_reference(net, size, size, cam, scenes.DEFAULT_FOV_Y, plume_tf, step, variant="h32x4_relu_alpha_noearly",
                                                  steps_variant="h32x4_relu_alpha_noearly_steps", density_range=rng)
    assert want[:, 3].max() > 0.3
    kwargs = dict(blend="alpha", early_out=False, density_range=rng)
    img, steps = parity.render_ours(volnet.dumps(net), "reference_exact", size, size, cam, plume_tf, step, **kwargs)
    _assert_identical(img, steps, want, want_steps, "alpha_noearly")
    for mode in ("tcgen05", "mma_sync"):
        img, steps = parity.render_ours(volnet.dumps(net), mode, size, size, cam, plume_tf, step, **kwargs)
        r = parity.compare(img, steps, want, want_steps)
        assert r["rgba_max_abs"] <= parity.RGBA_TOL and r["psnr_db"] >= parity.PSNR_MIN and r["step_count_mismatch_frac"] == 0.0, (mode, r)


def test_camera_batch_with_per_view_transfer_functions(ref, plume_tf):
    """B = 3 cameras in one call, one TF per view (the batch index of ImageEvaluatorSimpleKernel selects camera and TF,
    renderer_image_evaluator_simple.cuh:60-98): reference-exact mode bit-identical, fast modes within the image gate."""
    import parity
    from fvsrn_b200 import scenes, volnet
    kw, _, step = CASES["C3_relu_nerf"][:3]
    size = 256
    net = make_case_network(kw)
    cams = scenes.camera_sweep(3)
    tfs = np.concatenate([plume_tf, plume_tf, plume_tf], 0).copy()
    tfs[1, :, 3] *= 0.4; tfs[2, :, 3] *= 2.5; tfs[2, :, :3] = tfs[2, :, :3][:, ::-1]
    want, want_steps, _ = parity.render_reference(net, size, size, cams, scenes.DEFAULT_FOV_Y, tfs, step)
    assert want.shape[0] == 3 and not np.array_equal(want[0, :4], want[1, :4])
    img, steps = parity.render_ours(volnet.dumps(net), "reference_exact", size, size, cams, tfs, step)
    _assert_identical(img, steps, want, want_steps, "batch3")
    for mode in ("tcgen05", "mma_sync"):
        img, steps = parity.render_ours(volnet.dumps(net), mode, size, size, cams, tfs, step)
        r = parity.compare(img, steps, want, want_steps)
        assert r["rgba_max_abs"] <= parity.RGBA_TOL and r["psnr_db"] >= parity.PSNR_MIN, (mode, r)


def test_time_input_with_interpolated_keyframes(ref, plume_tf):
    """PASS_TIME_TO_NETWORK: the keyframe time is network input channel 3 and its fractional part blends the two keyframe
    grids (renderer_volume_tensorcores.cuh:590-604, 777-779): reference-exact mode bit-identical incl. step counts."""
    import parity
    from fvsrn_b200 import scenes, volnet
    net = scenes.make_network(**dict(CASES["C3_relu_nerf"][0], pass_time=True, time_keyframes=4))
    cam = scenes.camera_reference_frame(**scenes.DEFAULT_CAMERA)
    size, step, time = 512, 1 / 256, 1.37
    want, want_steps, _ = parity.render_reference(net, size, size, cam, scenes.DEFAULT_FOV_Y, plume_tf, step, variant="h32x4_relu_time",
                                                  steps_variant="h32x4_relu_time_steps", time=time)
    assert want[:, 3].max() > 0.3
    img, steps = parity.render_ours(volnet.dumps(net), "reference_exact", size, size, cam, plume_tf, step, time=time)
    _assert_identical(img, steps, want, want_steps, "time_input")


@pytest.mark.parametrize("name", ["direction", "direction_in_fourier", "latent_32_channels"])
def test_exact_mode_with_direction_input_and_two_latent_groups(ref, name, plume_tf):
    """USE_DIRECTION 1 / 2 and LATENT_GRID_CHANNELS_DIV16 2 (renderer_volume_tensorcores.cuh:784-804, 839-910): reference-exact
    mode bit-identical to the reference kernel built with that define, incl. step counts."""
    import parity
    from fvsrn_b200 import scenes, volnet
    from test_gpu_baseline_configs import INPUT_CASES
    kw, variant, time = INPUT_CASES[name]
    net = scenes.make_network(**kw)
    cam = scenes.camera_reference_frame(**scenes.DEFAULT_CAMERA)
    size, step = 512, 1 / 256
    want, want_steps, _ = parity.render_reference(net, size, size, cam, scenes.DEFAULT_FOV_Y, plume_tf, step, variant=variant,
                                                  steps_variant=variant + "_steps", time=time)
    assert want[:, 3].max() > 0.2
    img, steps = parity.render_ours(volnet.dumps(net), "reference_exact", size, size, cam, plume_tf, step)
    _assert_identical(img, steps, want, want_steps, name)


def test_identity_transfer_function(ref):
    """TransferFunctionIdentity (renderer_tf_identity.cuh:36-55: rgb = density * scaleEmission, absorption = density *
    scaleAbsorption * stepsize): reference-exact mode bit-identical, fast modes within the image gate."""
    import parity
    from fvsrn_b200 import scenes, volnet
    kw, _, step = CASES["C3_relu_nerf"][:3]
    size = 512
    net = make_case_network(kw)
    cam = scenes.camera_reference_frame(**scenes.DEFAULT_CAMERA)
    tf = np.array([30.0, 0.8], np.float32)      # scale of the absorption, of the emission
    want, want_steps, _ = parity.render_reference(net, size, size, cam, scenes.DEFAULT_FOV_Y, tf, step, tf_kind=2)
    assert 0.2 < want[:, 3].max() < 0.95
    img, steps = parity.render_ours(volnet.dumps(net), "reference_exact", size, size, cam, tf, step, tf_kind=2)
    _assert_identical(img, steps, want, want_steps, "identity_tf")
    for mode in ("tcgen05", "mma_sync"):
        img, steps = parity.render_ours(volnet.dumps(net), mode, size, size, cam, tf, step, tf_kind=2)
        r = parity.compare(img, steps, want, want_steps)
        assert r["rgba_max_abs"] <= parity.RGBA_TOL and r["psnr_db"] >= parity.PSNR_MIN, (mode, r)
