"""
Scene settings files (fv-srn_b200/settings.py; reference: ModuleRegistry::loadTree + each module's `load`, cited there): the
transfer-function assembly done at load time and the module walk. CPU only. The reference holds no test or golden vector for
these functions; they are checked against what they must represent - the merged piecewise TF is, at every density, the colour
lerp of the colour points and the absorption lerp of the opacity points - and, where /root/reference is present, against
every settings file the reference ships (each parses, or is refused for a module that is not on the path).
"""
import glob
import json
import os

import numpy as np
import pytest

from fvsrn_b200 import settings


def _lerp_points(pts, d):
    pts = sorted(pts, key=lambda p: p[0])
    xs = [p[0] for p in pts]
    if d <= xs[0]: return np.array(pts[0][1:], np.float64)
    if d >= xs[-1]: return np.array(pts[-1][1:], np.float64)
    i = max(k for k in range(len(xs)) if xs[k] <= d)
    f = (d - xs[i]) / (xs[i + 1] - xs[i])
    return (1 - f) * np.array(pts[i][1:], np.float64) + f * np.array(pts[i + 1][1:], np.float64)


def test_piecewise_tensor_small_example_by_hand():
    tf = settings.piecewise_tf_tensor([[0.25, 1, 0, 0], [0.75, 0, 0, 1]], [[0.0, 0.0], [0.5, 1.0], [1.0, 0.5]], 20.0)
    # colour gets sentinels at -1 and 2; opacity starts at 0 and ends at 1; merged by position until one list is exhausted
    # (the opacity list ends at 1.0, so the colour sentinel at 2 is never reached)
    pos = tf[0, :, 4].tolist()
    assert pos == [-1.0, 0.25, 0.5, 0.75, 1.0]
    np.testing.assert_allclose(tf[0, :, 3], [0.0, 10.0, 20.0, 15.0, 10.0], rtol=1e-6)
    hi = np.float32(1) - np.float32(settings.FLT_EPSILON)
    np.testing.assert_allclose(tf[0, 1, :3], [hi, 0, 0]); np.testing.assert_allclose(tf[0, 3, :3], [0, 0, hi])
    np.testing.assert_allclose(tf[0, 2, :3], [0.5, 0, 0.5], atol=1e-7)                         # colour interpolated at the opacity point
    assert tf.dtype == np.float32 and tf.shape == (1, 5, 5)


def test_piecewise_tensor_represents_both_point_lists():
    rng = np.random.default_rng(5)
    for trial in range(40):
        nc, no = rng.integers(1, 6), rng.integers(2, 7)
        cps = [[float(x)] + rng.uniform(0, 0.99, 3).tolist() for x in rng.uniform(0, 1, nc)]
        ops = [[float(x), float(rng.uniform(0.05, 1))] for x in rng.uniform(0, 1, no)]      # no zero-absorption region: nothing is purged
        if trial % 3 == 0: ops[0][0], ops[-1][0] = 0.0, 1.0
        if trial % 4 == 0: cps[0][0] = 0.0
        scale = float(rng.uniform(1, 50))
        tf = settings.piecewise_tf_tensor(cps, ops, scale)
        assert np.all(np.diff(tf[0, :, 4]) >= 0)
        # the first merged point takes the smaller of the two first positions with BOTH first values (computeTensor :203-207): when one
        # list starts at 0 and the other got its sentinel at -1, the first point of the former moves to -1 and the interval up to the
        # second merged point is a lerp from there - kept as in the reference, excluded from the representation property
        both_sentinels = min(p[0] for p in cps) > 0 and min(p[0] for p in ops) > 0
        for d in rng.uniform(0, 1, 50):
            if not both_sentinels and d < tf[0, 1, 4]:
                continue
            got = settings.sample_piecewise(tf, d)
            np.testing.assert_allclose(got[:3], _lerp_points(cps, d), atol=3e-5, err_msg=f"trial {trial} colour at {d}")
            np.testing.assert_allclose(got[3], _lerp_points(ops, d)[0] * scale, rtol=2e-4, atol=2e-5 * scale, err_msg=f"trial {trial} absorption at {d}")


def test_piecewise_purges_colour_points_inside_zero_absorption_regions():
    cps = [[0.05, 1, 0, 0], [0.1, 0, 1, 0], [0.15, 0, 0, 1], [0.8, 1, 1, 0]]
    ops = [[0.0, 0.0], [0.5, 0.0], [1.0, 1.0]]
    tf = settings.piecewise_tf_tensor(cps, ops, 1.0)
    assert tf.shape[1] < 2 + len(cps) + len(ops)
    for d in np.linspace(0, 0.5, 11):
        assert settings.sample_piecewise(tf, d)[3] == 0.0
    for d in (0.6, 0.8, 0.95):          # outside the purged region both lists are represented exactly
        got = settings.sample_piecewise(tf, d)
        np.testing.assert_allclose(got[3], _lerp_points(ops, d)[0], atol=1e-6)
    np.testing.assert_allclose(settings.sample_piecewise(tf, 0.8)[:3], [1 - settings.FLT_EPSILON, 1 - settings.FLT_EPSILON, 0], atol=1e-6)
    with pytest.raises(RuntimeError, match="deserialize"):
        settings.piecewise_tf_tensor([[0.1, 1, 0]], ops, 1.0)


def test_texture_tensor_and_paste():
    cps = [[0.0, 0, 0, 0], [0.5, 1, 0.5, 0], [1.0, 1, 1, 1]]
    plot = np.linspace(0, 1, 256).astype(np.float32).tolist()
    tex = settings.texture_tf_tensor(cps, plot, 7.0)
    assert tex.shape == (1, 256, 4) and tex.dtype == np.float32
    for i in (0, 17, 127, 128, 255):
        d = (i + 0.5) / 256
        np.testing.assert_allclose(tex[0, i, :3], _lerp_points(cps, d), atol=1e-6)
        np.testing.assert_allclose(tex[0, i, 3], 7.0 * plot[i], rtol=1e-6)
    with pytest.raises(RuntimeError, match="256"):
        settings.texture_tf_tensor(cps, plot[:100], 1.0)
    # doPaste: a piecewise TF resampled at the texel centres; absorption above the texture's scaling (10) raises the scaling
    tf = settings.piecewise_tf_tensor(cps, [[0.0, 0.0], [1.0, 1.0]], 25.0)
    pcps, pplot, scale = settings.paste_into_texture(lambda d: settings.sample_piecewise(tf, d))
    assert abs(scale - 25.0 * (255.5 / 256)) < 1e-4 and max(pplot) <= 1.0 + 1e-6
    tex2 = settings.texture_tf_tensor(pcps, pplot, scale)
    for i in (3, 100, 250):
        np.testing.assert_allclose(tex2[0, i], settings.sample_piecewise(tf, (i + 0.5) / 256), rtol=1e-5, atol=1e-6)


def _settings_dict(**over):
    s = {"version": 2, "root": "Simple",
         "ImageEvaluator": {"Simple": {"doublePrecision": False, "outputChannel": "ChannelColor", "samplesPerIterationLog2": 3, "selectedCamera": "Sphere",
                                       "selectedRayEvaluator": "DVR", "useTonemapping": False, "fixMaxExposure": True, "fixedMaxExposure": 2.5,
                                       "tonemappingShoulder": 0.5}},
         "RayEvaluation": {"DVR": {"maxDensity": 0.9, "minDensity": 0.1, "selectedBRDF": "Lambert", "selectedTF": "Piecewise", "selectedVolume": "Grid",
                                   "stepsize": 0.512, "stepsizeIsObjectSpace": True},
                           "Iso": {"isovalue": 0.5}},
         "camera": {"Sphere": {"center": [0.0, -0.15, 0.1], "distance": 1.5, "fovY": 0.7, "orientation": "Zp", "pitch": 0.3, "yaw": 2.1, "zoom": -1.0}},
         "brdf": {"Lambert": {"ambient": 0.2, "enablePhong": False, "enableMagnitudeScaling": False, "lightType": "Point", "lightPosition": [1, 2, 3]}},
         "blending": {"blending": {"blending": "Alpha"}},
         "tf": {"Piecewise": {"absorptionScaling": 25.0, "colorPoints": [[0.0, 0.1, 0.2, 0.3], [1.0, 0.9, 0.8, 0.7]], "opacityPoints": [[0.05, 0.0], [1.0, 0.9]]},
                "Identity": {"absorptionScaling": 3.0, "emissionScaling": 2.0},
                "Gaussian": {"absorptionScaling": 10.0, "points": []}},
         "volume": {"Grid": {"interpolation": "TRILINEAR", "mipmapLevel": 0, "source": "VOLUME", "volumePath": "../volumes/a.cvol"}}}
    for k, v in over.items():
        tag, name, key = k.split("__")
        s[tag][name][key] = v
    return s


def test_parse_settings_reads_what_the_modules_load():
    plan = settings.parse_settings(_settings_dict(), "/data/configs")
    assert plan["evaluator"]["channel"] == 3 and plan["evaluator"]["fixed_max_exposure"] == 2.5 and plan["evaluator"]["fix_max_exposure"]
    assert plan["camera"] == {"orientation": 4, "center": (0.0, -0.15, 0.1), "pitch": 0.3, "yaw": 2.1, "distance": 1.5, "fov_y": 0.7, "zoom": -1.0}
    assert plan["dvr"] == {"stepsize": 0.512 / 256, "min_density": 0.1, "max_density": 0.9, "early_out": True}     # old object-space step sizes: / 256
    assert plan["volume"]["path"] == "/data/volumes/a.cvol" and plan["volume"]["kind"] == "Grid"
    assert plan["tf"]["kind"] == "Piecewise" and plan["tf"]["absorption_scaling"] == 25.0
    assert plan["brdf"]["light_type"] == 0 and plan["brdf"]["light_position"] == (1.0, 2.0, 3.0) and plan["brdf"]["specular"] == 1.0    # default of ::load
    assert plan["file_blending"] == "Alpha"
    plan = settings.parse_settings(_settings_dict(RayEvaluation__DVR__selectedTF="Identity", RayEvaluation__DVR__stepsizeIsObjectSpace=False,
                                                  RayEvaluation__DVR__selectedVolume="SRN"), "/x")
    assert plan["tf"] == {"kind": "Identity", "absorption_scaling": 3.0, "emission_scaling": 2.0} and plan["dvr"]["stepsize"] == 0.512
    assert plan["volume"] == {"kind": "Network", "gradient_mode": 0, "finite_differences_stepsize": 1 / 256.0}
    for over, msg in ((dict(RayEvaluation__DVR__selectedTF="Gaussian"), "transfer function 'Gaussian'"),
                      (dict(ImageEvaluator__Simple__selectedRayEvaluator="Iso"), "ray evaluator 'Iso'"),
                      (dict(RayEvaluation__DVR__selectedVolume="Implicit"), "volume module 'Implicit'"),
                      (dict(ImageEvaluator__Simple__doublePrecision=True), "doublePrecision")):
        with pytest.raises(RuntimeError, match=msg):
            settings.parse_settings(_settings_dict(**over), "/x")
    bad = _settings_dict(); bad["root"] = "Neural Texture"
    with pytest.raises(RuntimeError, match="image evaluator"):
        settings.parse_settings(bad, "/x")
    with pytest.raises(RuntimeError, match="Unable to open file"):
        settings.read_settings("/nonexistent/settings.json")


def test_every_settings_file_of_the_reference_parses_or_is_refused_by_name(tmp_path):
    files = sorted(glob.glob("/root/reference/applications/**/config-files/*.json", recursive=True))
    if not files:
        pytest.skip("/root/reference is not present on this machine")
    loaded = refused = 0
    for f in files:
        with open(f) as fh:
            if "root" not in json.load(fh):
                continue            # not a renderer settings file
        try:
            plan = settings.read_settings(f)
        except RuntimeError as e:
            assert "not part of the B200 path" in str(e), (f, str(e))
            refused += 1
            continue
        loaded += 1
        if plan["tf"] and plan["tf"]["kind"] == "Piecewise":
            tf = settings.piecewise_tf_tensor(plan["tf"]["color_points"], plan["tf"]["opacity_points"], plan["tf"]["absorption_scaling"])
            assert np.isfinite(tf).all() and np.all(np.diff(tf[0, :, 4]) >= 0) and tf[0, :, 3].max() <= plan["tf"]["absorption_scaling"] * (1 + 1e-6)
        if plan["tf"] and plan["tf"]["kind"] == "Texture":
            assert np.isfinite(settings.texture_tf_tensor(plan["tf"]["color_points"], plan["tf"]["plot"], plan["tf"]["absorption_scaling"])).all()
    assert loaded >= 5 and loaded + refused >= 20, (loaded, refused)
