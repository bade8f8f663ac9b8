"""SURVEY f4 on the CPU: properties of the oracle's normals / shading restatement (oracle/fvsrn_oracle.c, test infrastructure) and
the host-side handling of the gradient output modes. The GPU kernels are compared with the reference's own kernels in
test_gpu_gradients.py."""
import numpy as np
import pytest

import oracle
from fvsrn_b200 import scenes, volnet

C3 = dict(hidden=32, num_layers=4, grid_res=32, seed=42, activation="ReLU", last_gain=6.0)
KW = dict(width=32, height=32, fov_y=scenes.DEFAULT_FOV_Y, stepsize=1 / 64)
PHONG = dict(flags=("magnitude", "phong"), light_type=1, light=(0.3, -0.5, -0.8), magnitude_scaling=2.0, ambient=0.3, specular=0.5,
             magnitude_center=1.0, magnitude_radius=0.8, specular_exponent=8)


def _cam():
    return scenes.camera_reference_frame(**scenes.DEFAULT_CAMERA)


def test_finite_difference_normals_leave_unshaded_colour_alone(plume_tf):
    net = scenes.make_network(**C3)
    plain, steps0, _ = oracle.render_dvr(net, camera=_cam(), tf=plume_tf, **KW)
    fd, steps1, _ = oracle.render_dvr(net, camera=_cam(), tf=plume_tf, fd_stepsize=1 / 64, **KW)
    assert np.array_equal(plain[:, :4], fd[:, :4]) and np.array_equal(steps0, steps1)
    assert np.all(plain[:, 4:7] == 0) and np.abs(fd[:, 4:7]).max() > 0.02
    # the normal channel is a sum of unit normals weighted like the colour: its length never exceeds alpha (squared, per sample)
    assert np.all(np.linalg.norm(fd[:, 4:7], axis=1) <= fd[:, 3] + 1e-5)


def test_predicted_normals_and_cubic_mode(plume_tf):
    lin = scenes.make_network(**dict(C3, output_mode="densitygrad:direct"))
    cub = scenes.make_network(**dict(C3, output_mode="densitygrad:cubic"))        # same weights, gradient outputs cubed
    a, _, _ = oracle.render_dvr(lin, camera=_cam(), tf=plume_tf, **KW)
    b, _, _ = oracle.render_dvr(cub, camera=_cam(), tf=plume_tf, **KW)
    assert np.array_equal(a[:, :4], b[:, :4])            # same density channel -> same colour
    assert np.abs(a[:, 4:7]).max() > 0.05 and not np.array_equal(a[:, 4:7], b[:, 4:7])
    assert oracle.evaluate(lin, np.random.default_rng(0).random((64, 3), np.float32)).shape == (64, 1)


def test_phong_shading_changes_colour_not_geometry(plume_tf):
    net = scenes.make_network(**dict(C3, output_mode="densitygrad:direct"))
    plain, s0, _ = oracle.render_dvr(net, camera=_cam(), tf=plume_tf, **KW)
    phong, s1, _ = oracle.render_dvr(net, camera=_cam(), tf=plume_tf, brdf=dict(PHONG, flags=("phong",)), **KW)
    assert np.array_equal(plain[:, 3], phong[:, 3]) and np.array_equal(s0, s1)     # Phong touches rgb only
    assert np.abs(plain[:, :3] - phong[:, :3]).max() > 1e-3
    mag, _, _ = oracle.render_dvr(net, camera=_cam(), tf=plume_tf, brdf=dict(PHONG, flags=("magnitude",)), **KW)
    assert np.all(mag[:, 3] <= plain[:, 3] + 1e-6) and (mag[:, 3] < plain[:, 3] - 1e-4).any()    # magnitude scaling only removes absorption


@pytest.mark.parametrize("mode", volnet.OUTPUT_MODES[4:])
def test_gradient_output_modes_round_trip(mode):
    net = scenes.make_network(**dict(C3, output_mode=mode))
    assert net.layers[-1].cout == (6 if "curvature" in mode else 4)
    back = volnet.loads(volnet.dumps(net))
    assert back.output_mode == mode and np.array_equal(back.layers[-1].weights, net.layers[-1].weights)


@pytest.mark.parametrize("kw", [dict(activation="ReLU", output_mode="density:direct"), dict(activation="SnakeAlt", output_mode="density"),
                                dict(activation="Sine:1.5", output_mode="density:direct"), dict(activation="Snake:2", output_mode="densitygrad")],
                         ids=["relu_direct", "snakealt_sigmoid", "sine_direct", "snake_densitygrad"])
def test_adjoint_restatement_is_the_gradient_of_the_network(kw):
    """GRADIENT_MODE 2 restated (fvsrn_oracle.c adjoint_h, following _evalNormalAdjoint :1205-1530) against what it must be: the
    derivative of the density output with respect to the unit-box position - central differences (step 2e-4) of the fp32
    evaluation of the same weights. Sample points at the centres of the trilinear cells, where the reference's quarter-voxel
    central difference through the latent grid IS the cell's slope. What remains is the fp16 arithmetic of the adjoint pass
    (and, for ReLU, samples next to a kink): measured cosine 0.9957 ... 0.9999, rms error 3 ... 11 % of the gradient's rms."""
    net = scenes.make_network(hidden=32, num_layers=4, grid_res=32, seed=42, last_gain=6.0, **kw)
    cells = np.random.default_rng(1).integers(2, 29, (200, 3))
    pos = ((cells + 1) / 32.0).astype(np.float32)
    g = oracle.adjoint_gradient(net, pos)
    eps = 2e-4
    fd = np.zeros_like(g)
    for a in range(3):
        e = np.zeros(3, np.float32); e[a] = eps
        fd[:, a] = (oracle.evaluate(net, pos + e, fp32=True)[:, 0].astype(np.float64) - oracle.evaluate(net, pos - e, fp32=True)[:, 0]) / (2 * eps)
    cos = np.sum(g * fd, 1) / (np.linalg.norm(g, axis=1) * np.linalg.norm(fd, axis=1) + 1e-12)
    rms = float(np.sqrt((fd ** 2).mean()))
    assert rms > 0.1 and cos.mean() >= 0.99 and float(np.sqrt(((g - fd) ** 2).mean())) <= 0.15 * rms, (cos.mean(), rms)


def test_adjoint_frame_shades_like_the_finite_difference_frame(plume_tf):
    net = scenes.make_network(**C3)
    plain, steps0, _ = oracle.render_dvr(net, camera=_cam(), tf=plume_tf, **KW)
    adj, steps1, _ = oracle.render_dvr(net, camera=_cam(), tf=plume_tf, adjoint=True, **KW)
    fd, _, _ = oracle.render_dvr(net, camera=_cam(), tf=plume_tf, fd_stepsize=1 / 256, **KW)
    assert np.array_equal(plain[:, :4], adj[:, :4]) and np.array_equal(steps0, steps1)      # normals without a BRDF leave the colour alone
    assert np.abs(adj[:, 4:7]).max() > 0.02
    # unit normals, blended like the colour: the two gradient modes agree on the blended normal (box size 1: same coordinates)
    assert np.abs(adj[:, 4:7] - fd[:, 4:7]).mean() <= 5e-3
    pa, _, _ = oracle.render_dvr(net, camera=_cam(), tf=plume_tf, adjoint=True, brdf=PHONG, **KW)
    pf, _, _ = oracle.render_dvr(net, camera=_cam(), tf=plume_tf, fd_stepsize=1 / 256, brdf=PHONG, **KW)
    assert np.abs(pa[:, :4] - pf[:, :4]).max() <= 2e-2 and np.abs(pa[:, :4] - pf[:, :4]).mean() <= 5e-4
