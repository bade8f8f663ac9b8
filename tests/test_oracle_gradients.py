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
