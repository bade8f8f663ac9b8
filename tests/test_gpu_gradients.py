"""
SURVEY f4 - normals and shading in the DVR frame, against the reference's own kernels (oracle/_ref, built from the unmodified
renderer/*.cuh with the matching defines): networks that predict the gradient next to the density (densitygrad, densitygrad:cubic),
finite-difference normals on a density network (GRADIENT_MODE 1, 6 more SRN evaluations per shaded step), BRDFLambert with
gradient-magnitude scaling and Phong shading (directional and point light), the normal channels 4..6 of the output tensor.
  * reference-exact kernel mode: all 8 channels bit-identical to the reference kernel;
  * fast modes (the mma.sync family renders every frame that needs normals, whichever fast mode is selected): RGBA within the
    north_star tolerance where the normals come from the network; for finite differences the fp16 noise of two evaluations is
    divided by 2h, so the gate is 3e-3 / 70 dB (measured 1.94e-3 / 85 dB; numbers in gpurun_out/parity_gradients.jsonl).
"""
import json
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
C3 = dict(hidden=32, num_layers=4, grid_res=32, seed=42, activation="ReLU")
PHONG_DIR = dict(flags=("magnitude", "phong"), light_type=1, light=(0.3, -0.5, -0.8), magnitude_scaling=2.0, ambient=0.3, specular=0.5,
                 magnitude_center=1.0, magnitude_radius=0.8, specular_exponent=8)
PHONG_POINT = dict(PHONG_DIR, light_type=0, light=(1.5, 1.0, -2.0))
# the cubed gradients of the random-init network are small (|g|^2 ~ 0.005, "gradientNorm" = 1/|g| ~ 10): scale the BRDF to that range
PHONG_DIR_CUBIC = dict(PHONG_DIR, magnitude_scaling=400.0, magnitude_center=8.0, magnitude_radius=6.0)
CASES = {
    # name: (network kwargs, reference variant, finite-difference step or None, brdf or None)
    "densitygrad": (dict(C3, output_mode="densitygrad", last_gain=6.0), "h32x4_relu_densgrad", None, None),
    "densitygrad_cubic_phong_directional": (dict(C3, output_mode="densitygrad:cubic", last_gain=6.0), "h32x4_relu_densgrad_cubic_phong", None, PHONG_DIR_CUBIC),
    "finite_differences_phong_point": (dict(C3, output_mode="density:direct", last_gain=6.0), "h32x4_relu_fd_phong", 1.0 / 128.0, PHONG_POINT),
}
# GRADIENT_MODE 2: the adjoint pass through the network (renderer_volume_tensorcores.cuh:1205-1530). ReLU / density:direct and
# SnakeAlt / density (sigmoid output: its adjoint runs in double, renderer_activations.cuh:204-209)
ADJOINT_CASES = {
    "adjoint_relu_phong_directional": (dict(C3, output_mode="density:direct", last_gain=6.0), "h32x4_relu_adjoint_phong", PHONG_DIR),
    "adjoint_snakealt_sigmoid_phong_point": (dict(C3, activation="SnakeAlt", output_mode="density", last_gain=6.0), "h32x4_snakealt_adjoint_phong", PHONG_POINT),
    # a network that also predicts a gradient: the adjoint differentiates its density output (sWeightsLast[2 * cin].x, :1256-1261)
    "adjoint_densitygrad_phong": (dict(C3, output_mode="densitygrad", last_gain=6.0), "h32x4_relu_densgrad_adjoint", PHONG_DIR),
    "adjoint_sine": (dict(C3, activation="Sine:1.5", last_gain=1.5, last_bias=-0.35), "h32x4_sine_adjoint", None),
    "adjoint_snake": (dict(C3, activation="Snake:2", last_gain=2.5, last_bias=-0.3), "h32x4_snake_adjoint", None),
    "adjoint_32_latent_channels": (dict(C3, grid_channels=32, last_gain=6.0), "h32x4_relu_grid32_adjoint", None),
    "adjoint_width_48": (dict(hidden=48, num_layers=4, grid_res=32, seed=42, activation="ReLU", fourier="gauss", last_gain=6.0, last_bias=0.05), "h48x4_relu_adjoint", None),
}
SIZE, STEP = 512, 1 / 256


@pytest.fixture(scope="module")
def ref():
    import torch
    assert torch.cuda.is_available()
    import ref_kernel
    if not ref_kernel.available():
        pytest.skip("oracle/_ref/*.ptx not built (needs /root/reference: python oracle/build_ref_ptx.py)")
    return ref_kernel


def _record(row):
    d = os.path.join(ROOT, "gpurun_out")
    if os.path.isdir(d):
        with open(os.path.join(d, "parity_gradients.jsonl"), "a") as f:
            f.write(json.dumps(row) + "\n")


def _frames(name, plume_tf, modes):
    import parity
    import fvsrn_b200  # noqa: F401
    from fvsrn_b200 import scenes, volnet
    kw, variant, fd, brdf = CASES[name]
    net = scenes.make_network(**kw)
    cam = scenes.camera_reference_frame(**scenes.DEFAULT_CAMERA)
    want, _, _ = parity.render_reference(net, SIZE, SIZE, cam, scenes.DEFAULT_FOV_Y, plume_tf, STEP, variant=variant,
                                         fd_stepsize=fd if fd else 0.01, brdf=brdf)
    data = volnet.dumps(net)
    got = {m: parity.render_ours(data, m, SIZE, SIZE, cam, plume_tf, STEP, fd_stepsize=fd, brdf=brdf)[0] for m in modes}
    return want, got


@pytest.mark.parametrize("name", list(CASES))
def test_reference_exact_mode_is_bit_identical_with_normals(ref, name, plume_tf):
    want, got = _frames(name, plume_tf, ["reference_exact"])
    img = got["reference_exact"]
    assert want[:, 3].max() > 0.1 and np.abs(want[:, 4:7]).max() > 0.02        # a frame with content and normals
    for c, label in enumerate(("r", "g", "b", "alpha", "normal x", "normal y", "normal z", "depth")):
        same = (img[:, c] == want[:, c]) | (np.isnan(img[:, c]) & np.isnan(want[:, c]))
        assert same.all(), (name, label, int((~same).sum()), float(np.nanmax(np.abs(img[:, c] - want[:, c]))))


@pytest.mark.parametrize("name", list(CASES))
def test_fast_modes_with_normals(ref, name, plume_tf):
    import parity
    want, got = _frames(name, plume_tf, ["mma_sync", "tcgen05"])
    fd = CASES[name][2] is not None
    for mode, img in got.items():
        d = np.abs(img[:, :4] - want[:, :4])
        dn = np.abs(img[:, 4:7] - want[:, 4:7])
        row = dict(config=name, kernel_mode=mode, size=SIZE, rgba_max_abs=float(d.max()), psnr_db=float(parity.psnr(img[:, :4], want[:, :4])),
                   normal_max_abs=float(dn.max()), normal_mean_abs=float(dn.mean()), frac_rgba_over_2e3=float((d.max(1) > 2e-3).mean()))
        _record(row)
        if not fd:
            assert row["rgba_max_abs"] <= parity.RGBA_TOL and row["psnr_db"] >= parity.PSNR_MIN, row
            assert row["normal_max_abs"] <= 1e-2, row
        else:
            # central differences of two fp16-accumulated evaluations, divided by 2h = 1/64: the normal of a sample moves with
            # the last bits of the network output. Measured on B200: 1.94e-3 RGBA max-abs, 85 dB, normals 1.2e-3 max-abs
            assert row["psnr_db"] >= 70.0 and row["rgba_max_abs"] <= 3e-3, row
            assert row["normal_max_abs"] <= 1e-2 and row["normal_mean_abs"] <= 1e-4, row


def _adjoint_frames(name, plume_tf, modes, size=SIZE):
    import parity
    import fvsrn_b200  # noqa: F401
    from fvsrn_b200 import scenes, volnet
    kw, variant, brdf = ADJOINT_CASES[name]
    net = scenes.make_network(**kw)
    cam = scenes.camera_reference_frame(**scenes.DEFAULT_CAMERA)
    want, _, _ = parity.render_reference(net, size, size, cam, scenes.DEFAULT_FOV_Y, plume_tf, STEP, variant=variant, brdf=brdf)
    data = volnet.dumps(net)
    got = {m: parity.render_ours(data, m, size, size, cam, plume_tf, STEP, brdf=brdf, adjoint=(m != "finite_differences"),
                                 fd_stepsize=(1.0 / 256.0 if m == "finite_differences" else None))[0]
           for m in modes}
    return want, got


@pytest.mark.parametrize("name", list(ADJOINT_CASES))
def test_adjoint_gradient_mode_is_bit_identical_to_the_reference_kernel(ref, name, plume_tf):
    """Reference-exact kernel mode: all 8 channels of a frame in adjoint gradient mode bit-identical to the reference kernel built
    with GRADIENT_MODE 2."""
    want, got = _adjoint_frames(name, plume_tf, ["reference_exact"])
    assert want[:, 3].max() > 0.1 and np.abs(want[:, 4:7]).max() > 0.02
    for mode, img in got.items():
        for c, label in enumerate(("r", "g", "b", "alpha", "normal x", "normal y", "normal z", "depth")):
            same = (img[:, c] == want[:, c]) | (np.isnan(img[:, c]) & np.isnan(want[:, c]))
            assert same.all(), (name, mode, label, int((~same).sum()), float(np.nanmax(np.abs(img[:, c] - want[:, c]))))


@pytest.mark.parametrize("name", list(ADJOINT_CASES))
def test_fast_modes_with_adjoint_normals(ref, name, plume_tf):
    """The fast kernel modes run the adjoint pass in their own arithmetic (mma.sync family: SFU sin / cos, exact trilinear weights,
    fp32 Fourier adjoint). The adjoint is a chain of fp16 products through every layer (ReLU masks flip on the last bit of a
    pre-activation), so two fp16 implementations differ by the noise of the method itself: on these networks the reference's own
    adjoint is 1 ... 11 % (rms) away from the fp32 gradient (tests/test_oracle_gradients.py). The gate is therefore the fp32 frame -
    central differences of the fp32 network, CPU oracle - and the criterion "at least as close to it as the reference kernel is"
    (factor 1.25 + 1e-4 on the mean normal error, PSNR of the shaded colour within 1 dB; measured: equal to 3 digits on all seven
    networks), next to bounds on the means against the reference frame."""
    import oracle
    import parity
    import fvsrn_b200  # noqa: F401
    from fvsrn_b200 import scenes, volnet
    kw, variant, brdf = ADJOINT_CASES[name]
    net = scenes.make_network(**kw)
    cam = scenes.camera_reference_frame(**scenes.DEFAULT_CAMERA)
    W, step = 96, 1 / 96
    truth, _, _ = oracle.render_dvr(net, width=W, height=W, camera=cam, fov_y=scenes.DEFAULT_FOV_Y, stepsize=step, tf=plume_tf, fp32=True,
                                    fd_stepsize=1e-3, brdf=brdf)
    want, _, _ = parity.render_reference(net, W, W, cam, scenes.DEFAULT_FOV_Y, plume_tf, step, variant=variant, brdf=brdf)
    data = volnet.dumps(net)
    assert np.abs(truth[:, 4:7]).max() > 0.02
    err_ref = float(np.abs(want[:, 4:7] - truth[:, 4:7]).mean())
    psnr_ref = float(parity.psnr(want[:, :4], truth[:, :4]))
    for mode in ("tcgen05", "mma_sync"):
        img = parity.render_ours(data, mode, W, W, cam, plume_tf, step, brdf=brdf, adjoint=True)[0]
        err = float(np.abs(img[:, 4:7] - truth[:, 4:7]).mean())
        row = dict(config=name, kernel_mode=mode, size=W, normal_mean_abs_vs_fp32=err, reference_normal_mean_abs_vs_fp32=err_ref,
                   psnr_db_vs_fp32=float(parity.psnr(img[:, :4], truth[:, :4])), reference_psnr_db_vs_fp32=psnr_ref,
                   rgba_max_abs_vs_reference=float(np.abs(img[:, :4] - want[:, :4]).max()), normal_max_abs_vs_reference=float(np.abs(img[:, 4:7] - want[:, 4:7]).max()),
                   psnr_db_vs_reference=float(parity.psnr(img[:, :4], want[:, :4])), normal_mean_abs_vs_reference=float(np.abs(img[:, 4:7] - want[:, 4:7]).mean()))
        _record(row)
        assert err <= 1.25 * err_ref + 1e-4 and row["psnr_db_vs_fp32"] >= psnr_ref - 1.0, row
        # against the reference frame itself: means, not maxima - where the gradient is nearly zero its direction is noise in both
        # implementations (measured maxima: 0.24 on a normal component, 0.08 on a Phong-shaded colour, in isolated pixels)
        assert row["psnr_db_vs_reference"] >= 50.0 and row["normal_mean_abs_vs_reference"] <= 5e-3, row


def test_adjoint_through_time_interpolated_keyframes(ref, plume_tf):
    """Between two keyframes the adjoint differentiates the time-interpolated texel (lerp of both grids' central differences,
    :662-676): bit-identical to the reference kernel at time 1.37 of a 4-keyframe network."""
    import parity
    import fvsrn_b200  # noqa: F401
    from fvsrn_b200 import scenes, volnet
    net = scenes.make_network(**dict(C3, output_mode="density:direct", last_gain=6.0, time_keyframes=4))
    cam = scenes.camera_reference_frame(**scenes.DEFAULT_CAMERA)
    want, _, _ = parity.render_reference(net, 256, 256, cam, scenes.DEFAULT_FOV_Y, plume_tf, STEP, variant="h32x4_relu_adjoint_phong", brdf=PHONG_DIR, time=1.37)
    img = parity.render_ours(volnet.dumps(net), "reference_exact", 256, 256, cam, plume_tf, STEP, brdf=PHONG_DIR, adjoint=True, time=1.37)[0]
    assert want[:, 3].max() > 0.1 and np.abs(want[:, 4:7]).max() > 0.02
    assert np.array_equal(np.nan_to_num(img, nan=-7.0), np.nan_to_num(want, nan=-7.0))


def test_adjoint_frame_against_the_cpu_oracle(plume_tf):
    """The adjoint frame of the product path against the CPU restatement of the reference's adjoint (oracle/fvsrn_oracle.c adjoint_h)
    - no reference build needed (this is the check __graft_entry__.smoke() style boxes can run). The oracle's libm cos / sin and
    exact central-difference samples differ from the kernel's hcos / hsin and fp32-texture fetches in the last fp16 bits, and the
    Phong term (exponent 8) of a 1/64 step carries a whole sample's weight: RGBA within 5e-3 (measured 2.9e-3), blended normals
    within 2e-2 max-abs / 5e-4 mean-abs (measured 4.7e-3 / 1.2e-5). The bit-exact pin of this mode is the reference-kernel test above."""
    import oracle
    import parity
    import fvsrn_b200  # noqa: F401
    from fvsrn_b200 import scenes, volnet
    net = scenes.make_network(**dict(C3, output_mode="density:direct", last_gain=6.0))
    cam = scenes.camera_reference_frame(**scenes.DEFAULT_CAMERA)
    W = 64
    want, _, _ = oracle.render_dvr(net, width=W, height=W, camera=cam, fov_y=scenes.DEFAULT_FOV_Y, stepsize=1 / 64, tf=plume_tf, adjoint=True, brdf=PHONG_DIR,
                                   tex_filter=True)
    img = parity.render_ours(volnet.dumps(net), "reference_exact", W, W, cam, plume_tf, 1 / 64, brdf=PHONG_DIR, adjoint=True)[0]
    d = np.abs(img[:, :4] - want[:, :4]); dn = np.abs(img[:, 4:7] - want[:, 4:7])
    row = dict(config="adjoint_vs_cpu_oracle", rgba_max_abs=float(d.max()), normal_max_abs=float(dn.max()), normal_mean_abs=float(dn.mean()))
    _record(row)
    assert np.abs(want[:, 4:7]).max() > 0.02
    assert row["rgba_max_abs"] <= 5e-3 and row["normal_max_abs"] <= 2e-2 and row["normal_mean_abs"] <= 5e-4, row


def test_adjoint_normals_agree_with_finite_differences(plume_tf):
    """The two ways to the same gradient (no reference needed): the adjoint pass and central differences of the same network
    shade the frame alike - PSNR of the shaded colour >= 35 dB, mean normal difference small. (They are not equal: fp16 noise in
    the differences, the latent grid differentiated with a coarse step in the adjoint, and the adjoint is the gradient in unit-box
    coordinates.)"""
    import parity
    import fvsrn_b200  # noqa: F401
    from fvsrn_b200 import scenes, volnet
    kw, _, brdf = ADJOINT_CASES["adjoint_relu_phong_directional"]
    net = scenes.make_network(**kw)
    cam = scenes.camera_reference_frame(**scenes.DEFAULT_CAMERA)
    data = volnet.dumps(net)
    a = parity.render_ours(data, "mma_sync", 256, 256, cam, plume_tf, STEP, brdf=brdf, adjoint=True)[0]
    f = parity.render_ours(data, "mma_sync", 256, 256, cam, plume_tf, STEP, brdf=brdf, fd_stepsize=1.0 / 256.0)[0]
    assert np.abs(a[:, 4:7]).max() > 0.02
    row = dict(config="adjoint_vs_fd", psnr_db=float(parity.psnr(a[:, :4], f[:, :4])), normal_mean_abs=float(np.abs(a[:, 4:7] - f[:, 4:7]).mean()),
               normal_max_abs=float(np.abs(a[:, 4:7] - f[:, 4:7]).max()))
    _record(row)
    assert row["psnr_db"] >= 35.0 and row["normal_mean_abs"] <= 2e-2, row


def test_gradient_mode_errors(plume_tf):
    """colour networks have no normals, adjoint mode refuses what the reference's adjoint does not define, BRDF flags are validated."""
    import torch
    import fvsrn_b200  # noqa: F401
    from fvsrn_b200 import pyrenderer as pr, scenes, volnet
    sn = pr.SceneNetwork.from_bytes(volnet.dumps(scenes.make_network(**dict(C3, output_mode="rgbo"))))
    vol = pr.VolumeInterpolationNetwork(); vol.set_network(sn)
    ie = pr.ImageEvaluatorSimple(); ie.volume = vol
    ie.camera.set_parameters(torch.from_numpy(scenes.camera_reference_frame(**scenes.DEFAULT_CAMERA)))
    vol.gradient_mode = pr.VolumeInterpolationNetwork.FINITE_DIFFERENCES
    with pytest.raises(RuntimeError, match="density network"):
        ie.render(64, 64)
    vol.gradient_mode = pr.VolumeInterpolationNetwork.ADJOINT_METHOD
    with pytest.raises(RuntimeError, match="density network"):
        ie.render(64, 64)
    sd = pr.SceneNetwork.from_bytes(volnet.dumps(scenes.make_network(**dict(C3, has_direction=True))))
    vd = pr.VolumeInterpolationNetwork(); vd.set_network(sd); vd.gradient_mode = pr.VolumeInterpolationNetwork.ADJOINT_METHOD
    ie.volume = vd
    ie.ray_evaluator.tf.tensor.value = torch.from_numpy(plume_tf).cuda()
    with pytest.raises(RuntimeError, match="direction input"):
        ie.render(64, 64)
    ie.volume = vol
    vol.gradient_mode = pr.VolumeInterpolationNetwork.OFF_OR_DIRECT
    ie.ray_evaluator.brdf.enable_phong = True
    with pytest.raises(RuntimeError, match="normals"):
        ie.render(64, 64)


def test_evaluate_returns_density_for_gradient_networks():
    """IVolumeInterpolation::evaluate of a densitygrad network is its density channel (1 output, sigmoid applied)."""
    import torch
    import fvsrn_b200  # noqa: F401
    import oracle
    from fvsrn_b200 import pyrenderer as pr, scenes, volnet
    net = scenes.make_network(**dict(C3, output_mode="densitygrad", last_gain=6.0))
    sn = pr.SceneNetwork.from_bytes(volnet.dumps(net))
    vol = pr.VolumeInterpolationNetwork(); vol.set_network(sn)
    pos = torch.rand((4096, 3), device="cuda")
    got = vol.evaluate(pos).cpu().numpy()
    assert got.shape == (4096, 1)
    want = oracle.evaluate(net, pos.cpu().numpy())
    assert np.abs(got[:, 0] - np.asarray(want).reshape(len(got), -1)[:, 0]).max() <= 2e-3
