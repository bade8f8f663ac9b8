"""
pyrenderer.load_from_json (renderer/module_registry.cpp:75-99) end to end on the GPU: a settings file in the reference's schema
plus the .cvol it names -> an ImageEvaluatorSimple that renders the same frame, bit for bit, as an evaluator configured by hand
with the values the reference's `load` functions would read; the piecewise TF assembled at load time evaluates on the GPU
(fvsrn_evaluate_tf) as its CPU copy does (TransferFunctionPiecewiseLinear::evaluate); convert_to_texture_tf resamples it at the
256 texel centres; an SRN volume comes back empty and renders once a network is set.
"""
import json
import os

import numpy as np
import pytest

from test_gpu_grid import blob_volume
from test_settings import _settings_dict

pytestmark = pytest.mark.gpu


@pytest.fixture()
def scene_dir(tmp_path):
    import fvsrn_b200  # noqa: F401
    from fvsrn_b200 import pyrenderer as pr
    (tmp_path / "volumes").mkdir(); (tmp_path / "configs").mkdir()
    vol = blob_volume(40)
    pr.Volume.from_numpy(vol).save(str(tmp_path / "volumes" / "a.cvol"))
    return tmp_path, vol


def _write(path, settings_dict):
    with open(path, "w") as f:
        json.dump(settings_dict, f)
    return str(path)


def test_load_from_json_renders_like_a_hand_configured_evaluator(scene_dir):
    import torch
    from fvsrn_b200 import pyrenderer as pr, scenes, settings
    root, vol = scene_dir
    s = _settings_dict(RayEvaluation__DVR__stepsize=1.0, RayEvaluation__DVR__minDensity=0.05, RayEvaluation__DVR__maxDensity=0.95)
    s["tf"]["Piecewise"] = {"absorptionScaling": 40.0, "colorPoints": [[0.1, 0.9, 0.2, 0.1], [0.5, 0.2, 0.8, 0.3], [0.9, 0.1, 0.3, 0.9]],
                            "opacityPoints": [[0.15, 0.0], [0.4, 0.5], [0.7, 0.2], [0.95, 1.0]]}
    s["camera"]["Sphere"].update(center=[0.0, 0.0, 0.0], distance=2.0, orientation="Ym", pitch=0.4, yaw=0.7)
    ie = pr.load_from_json(_write(root / "configs" / "scene.json", s))
    assert isinstance(ie.volume, pr.VolumeInterpolationGrid) and ie.volume.volume() is not None and ie.volume.mipmap_level() == 0
    assert abs(ie.ray_evaluator.stepsize - 1 / 256) < 1e-12 and ie.fix_max_exposure and ie.fixed_max_exposure == 2.5
    assert tuple(ie.camera.pitchYawDistance.value) == (0.4, 0.7, 2.0) and ie.camera.orientation == pr.CameraOnASphere.Ym
    assert ie.ray_evaluator.blending.blendMode == pr.Blending.BeerLambert          # the file says Alpha: not part of the loaded tree
    W, H = 160, 120
    got = ie.render(W, H).cpu().numpy()
    assert got[:, 3].max() > 0.3

    hand = pr.ImageEvaluatorSimple()
    g = pr.VolumeInterpolationGrid(); g.setSource(pr.Volume.from_numpy(vol)); hand.volume = g
    hand.camera.set_parameters(torch.from_numpy(scenes.camera_reference_frame(0.4, 0.7, 2.0, (0, 0, 0), 3)))
    hand.camera.fovY_radians = 0.7
    tf = settings.piecewise_tf_tensor(s["tf"]["Piecewise"]["colorPoints"], s["tf"]["Piecewise"]["opacityPoints"], 40.0)
    hand.ray_evaluator.tf.tensor.value = torch.from_numpy(tf).cuda()
    hand.ray_evaluator.stepsize, hand.ray_evaluator.min_density, hand.ray_evaluator.max_density = 1 / 256, 0.05, 0.95
    want = hand.render(W, H).cpu().numpy()
    assert np.array_equal(np.nan_to_num(got, nan=-7), np.nan_to_num(want, nan=-7))

    # the TF on the GPU against its CPU copy
    d = torch.linspace(0, 1, 513, device="cuda").reshape(-1, 1)
    gpu = ie.ray_evaluator.tf.evaluate(d, 0.0, 1.0).cpu().numpy()
    cpu = np.stack([np.array(tuple(ie.ray_evaluator.tf.evaluate_scalar(float(x))), np.float32) for x in d.cpu().numpy()[:, 0]])
    np.testing.assert_allclose(gpu, cpu, rtol=2e-5, atol=2e-6)
    assert ie.ray_evaluator.tf.get_max_absorption() == 40.0

    # convert_to_texture_tf: 256 texels of the same function; the frame moves by the resampling error only
    ie.ray_evaluator.convert_to_texture_tf()
    assert isinstance(ie.ray_evaluator.tf, pr.TransferFunctionTexture) and ie.ray_evaluator.tf.tensor.value.shape == (1, 256, 4)
    tex = ie.ray_evaluator.tf.tensor.value.cpu().numpy()[0]
    for i in (5, 80, 200, 250):
        np.testing.assert_allclose(tex[i], settings.sample_piecewise(tf, (i + 0.5) / 256), rtol=1e-5, atol=1e-6)
    again = ie.render(W, H).cpu().numpy()
    assert np.abs(again[:, :4] - got[:, :4]).max() < 2e-2 and np.abs(again[:, :4] - got[:, :4]).mean() < 5e-4


def test_load_from_json_with_texture_tf_identity_tf_and_srn_volume(scene_dir, plume_tf):
    import torch
    from fvsrn_b200 import pyrenderer as pr, scenes, volnet
    root, vol = scene_dir
    plot = (np.linspace(0, 1, 256) ** 2).astype(np.float32).tolist()
    s = _settings_dict(RayEvaluation__DVR__selectedTF="Texture", RayEvaluation__DVR__stepsizeIsObjectSpace=False, RayEvaluation__DVR__stepsize=1 / 128)
    s["tf"]["Texture"] = {"absorptionScaling": 30.0, "colorPoints": [[0.0, 1.0, 0.0, 0.0], [1.0, 1.0, 1.0, 1.0]], "opacityPoints": plot}
    ie = pr.load_from_json(_write(root / "configs" / "tex.json", s))
    assert isinstance(ie.ray_evaluator.tf, pr.TransferFunctionTexture)
    t = ie.ray_evaluator.tf.tensor.value.cpu().numpy()
    np.testing.assert_allclose(t[0, :, 3], 30.0 * np.asarray(plot), rtol=1e-6)
    assert ie.render(96, 64).cpu().numpy()[:, 3].max() > 0.2

    s = _settings_dict(RayEvaluation__DVR__selectedTF="Identity", RayEvaluation__DVR__selectedVolume="SRN", RayEvaluation__DVR__stepsizeIsObjectSpace=False,
                       RayEvaluation__DVR__stepsize=1 / 128, RayEvaluation__DVR__minDensity=0.0, RayEvaluation__DVR__maxDensity=1.0)
    s["volume"]["SRN"] = {"gradientMode": "FINITE_DIFFERENCES", "finiteDifferencesStepsize": 0.01}
    ie = pr.load_from_json(_write(root / "configs" / "srn.json", s))
    assert isinstance(ie.volume, pr.VolumeInterpolationNetwork) and ie.volume.gradient_mode == pr.VolumeInterpolationNetwork.FINITE_DIFFERENCES
    assert ie.volume.finite_differences_stepsize == 0.01 and isinstance(ie.ray_evaluator.tf, pr.TransferFunctionIdentity)
    with pytest.raises(RuntimeError, match="[Nn]o network"):
        ie.render(64, 64)
    net = scenes.make_network(hidden=32, num_layers=4, grid_res=16, seed=3, activation="ReLU", last_gain=6.0)
    ie.volume.gradient_mode = pr.VolumeInterpolationNetwork.OFF_OR_DIRECT
    ie.volume.set_network(pr.SceneNetwork.from_bytes(volnet.dumps(net)))
    ie.camera.pitchYawDistance.value = pr.double3(0.3, 0.5, 2.0)
    img = ie.render(96, 96).cpu().numpy()
    assert img[:, 3].max() > 0.2

    # a missing volume file: reported, the evaluator loads, rendering says what is missing (as in the reference)
    s = _settings_dict(volume__Grid__volumePath="../volumes/missing.cvol")
    ie = pr.load_from_json(_write(root / "configs" / "missing.json", s))
    with pytest.raises(RuntimeError, match="No volume specified"):
        ie.render(32, 32)
    with pytest.raises(RuntimeError, match="not part of the B200 path"):
        pr.load_from_json(_write(root / "configs" / "gauss.json", _settings_dict(RayEvaluation__DVR__selectedTF="Gaussian")))


def test_refine_folds_frames_into_a_running_mean(scene_dir):
    """IImageEvaluator.refine (image_evaluator_simple.cpp:288-356): n-th refinement = previous + (frame - previous) / n; DVR frames are
    deterministic, so the mean stays the first frame (to fp32 rounding) and the counter restarts with render()."""
    import torch
    from fvsrn_b200 import pyrenderer as pr
    root, vol = scene_dir
    ie = pr.load_from_json(_write(root / "configs" / "scene.json", _settings_dict(RayEvaluation__DVR__stepsize=1.0)))
    first = ie.render(96, 64)
    assert not ie.is_iterative_refining() and ie.compute_batch_count() == 1 and ie.spp_log2 == 3
    prev = torch.nan_to_num(first, nan=0.0) * 0.5           # a different "previous" shows the weights: 1/1, then 1/2, 1/3
    r1 = ie.refine(96, 64, prev)
    torch.testing.assert_close(torch.nan_to_num(r1, nan=0.0), prev + (torch.nan_to_num(first, nan=0.0) - prev), rtol=0, atol=1e-6)
    r2 = ie.refine(96, 64, prev)
    f = torch.nan_to_num(first, nan=0.0)
    torch.testing.assert_close(torch.nan_to_num(r2, nan=0.0), prev + (f - prev) * 0.5, rtol=0, atol=1e-6)
    ie.render(96, 64)
    r3 = ie.refine(96, 64, prev)
    torch.testing.assert_close(torch.nan_to_num(r3, nan=0.0), f, rtol=0, atol=1e-6)
    with pytest.raises(RuntimeError, match="previous"):
        ie.refine(96, 64, prev[:, :4])
    cams = torch.from_numpy(np.concatenate([np.asarray(ie.camera.get_parameters().cpu())] * 3, 0)).cuda()
    ie.camera.set_parameters(cams)
    assert ie.compute_batch_count() == 3
    ie.ray_evaluator.tf.tensor.value = torch.cat([ie.ray_evaluator.tf.tensor.value] * 2, 0)
    with pytest.raises(RuntimeError, match="Batch counts don't match"):
        ie.compute_batch_count()
