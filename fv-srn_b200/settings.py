"""
The reference's scene settings files (applications/config-files/*.json, written by ModuleRegistry::saveAll) as the DVR path reads
them: `pyrenderer.load_from_json` (renderer/module_registry.cpp:75-99 -> loadTree :288-305) walks from the selected image
evaluator through its selected camera / ray evaluator / volume / transfer function / BRDF and calls each module's `load`.
This file is the host side of that walk - pure NumPy, no device - and of the transfer-function assembly the modules do at
load time:
  * piecewise TF: colour points + opacity points -> the merged (1,R,5) tensor r,g,b,absorption,position
    (TransferFunctionPiecewiseLinear::computeTensor, transfer_function_piecewise.cpp:166-279);
  * texture TF: colour points + 256 opacity values -> (1,256,4) (TransferFunctionTexture::computeTexture, transfer_function_texture.cpp:347-362,
    TFPartPiecewiseColor::getAsTexture, transfer_function.cpp:526-552);
  * piecewise -> texture conversion (RayEvaluationSteppingDvr::convertToTextureTF -> TransferFunctionTexture::doPaste,
    ray_evaluation_stepping.cpp:757-770, transfer_function_texture.cpp:412-437).
Modules that are not on the B200 path (iso-surface / Monte-Carlo ray evaluators, Gaussian and 2-D TFs, implicit volumes, the
neural-texture evaluator) raise RuntimeError when the file selects them; unselected ones are skipped as in the reference.
"""
import json
import math
import os
from typing import Dict, List, Sequence

import numpy as np

FLT_EPSILON = float(np.finfo(np.float32).eps)
TEXTURE_RESOLUTION = 256              # TransferFunctionTexture::RESOLUTION, transfer_function.h:219


# ------------------------------------------------------------------------------------------------ transfer functions
def _sorted_points(points: Sequence[Sequence[float]], width: int, what: str) -> List[List[float]]:
    """from_json of the point types (transfer_function.h:403-433) + sortPoints (stable order is not promised by std::sort; ties keep
    file order here)."""
    pts = []
    for p in points:
        if not isinstance(p, (list, tuple)) or len(p) != width:
            raise RuntimeError(f"Unable to deserialize {p} into a {what} point")
        pts.append([float(v) for v in p])
    if not pts:
        raise RuntimeError(f"a transfer function needs at least one {what} point")
    return sorted(pts, key=lambda p: p[0])


def piecewise_tf_tensor(color_points, opacity_points, absorption_scaling: float) -> np.ndarray:
    """colorPoints [[pos, r, g, b], ...] + opacityPoints [[pos, absorption], ...] -> (1,R,5) float32 rows r,g,b,absorption,pos.
    Arithmetic in double like the reference, stored as float."""
    cp = [(p[0], np.array(p[1:4], np.float64)) for p in _sorted_points(color_points, 4, "TFPartPiecewiseColor")]
    op = [(p[0], p[1]) for p in _sorted_points(opacity_points, 2, "TfPartPiecewiseOpacity")]
    # sentinels outside [0, 1] (the reference: "better -1 and 2 as this is better for the iso-intersections")
    if cp[0][0] > 0: cp.insert(0, (-1.0, cp[0][1]))
    if op[0][0] > 0: op.insert(0, (-1.0, op[0][1]))
    if cp[-1][0] < 1: cp.append((2.0, cp[-1][1]))
    if op[-1][0] < 1: op.append((2.0, op[-1][1]))
    pts = [[min(cp[0][0], op[0][0]), cp[0][1], op[0][1]]]
    io = ic = 0
    with np.errstate(divide="ignore", invalid="ignore"):
        while io < len(op) - 1 and ic < len(cp) - 1:
            if op[io + 1][0] < cp[ic + 1][0]:       # next point is an opacity point: colour interpolated at its position
                f = np.float64(op[io + 1][0] - cp[ic][0]) / np.float64(cp[ic + 1][0] - cp[ic][0])
                pts.append([op[io + 1][0], cp[ic][1] + f * (cp[ic + 1][1] - cp[ic][1]), op[io + 1][1]])
                io += 1
            else:                                   # next point is a colour point: absorption interpolated
                f = np.float64(cp[ic + 1][0] - op[io][0]) / np.float64(op[io + 1][0] - op[io][0])
                pts.append([cp[ic + 1][0], cp[ic + 1][1], float(op[io][1] + f * (op[io + 1][1] - op[io][1]))])
                ic += 1
    # drop colour points inside zero-absorption regions and points closer than EPS (purgeZeroOpacityRegions)
    eps = float(np.float32(1e-7))
    i = 0
    while i < len(pts) - 2:
        if (pts[i][2] < eps and pts[i + 1][2] < eps and pts[i + 2][2] < eps) or (pts[i + 1][0] - pts[i][0] < eps):
            del pts[i + 1]
        else:
            i += 1
    out = np.empty((1, len(pts), 5), np.float32)
    hi = float(np.float32(1.0) - np.float32(FLT_EPSILON))       # clamp(rgb, 0.0f, 1.0f - FLT_EPSILON): float bounds
    for k, (pos, rgb, a) in enumerate(pts):
        out[0, k, :3] = np.clip(rgb, 0.0, hi)
        out[0, k, 3] = min(max(a, 0.0), 1.0) * absorption_scaling
        out[0, k, 4] = pos
    return out


def sample_piecewise(tf: np.ndarray, density: float) -> np.ndarray:
    """kernel::TransferFunctionPiecewise::sampleTF<float> on the CPU tensor (renderer_tf_piecewise.cuh:27-50), as
    TransferFunctionPiecewiseLinear::evaluate(double) calls it: float arithmetic, no clamp of the density to [0, 1]."""
    t = tf[0] if tf.ndim == 3 else tf
    d = np.float32(density)
    R = t.shape[0]
    i = 0
    while i < R - 2 and not (t[i + 1, 4] > d):
        i += 1
    p0, p1 = t[i, 4], t[i + 1, 4]
    d = min(max(d, p0), p1)
    frac = np.float32(d - p0) / np.float32(p1 - p0)
    return (t[i, :4] + frac * (t[i + 1, :4] - t[i, :4])).astype(np.float32)


def color_texture(color_points, resolution: int) -> np.ndarray:
    """TFPartPiecewiseColor::getAsTexture: (resolution,3) double; texel centres, float positions and fraction."""
    cp = _sorted_points(color_points, 4, "TFPartPiecewiseColor")
    n = len(cp)
    out = np.empty((resolution, 3), np.float64)
    with np.errstate(divide="ignore", invalid="ignore"):
        for i in range(resolution):
            density = np.float32(np.float32(i + 0.5) / np.float32(resolution))
            idx = 0
            while idx < n - 2 and not (cp[idx + 1][0] > float(density)):      # double position against the float density
                idx += 1
            nxt = min(idx + 1, n - 1)                       # a single colour point: the reference reads past the end (UB); use it twice
            lo, hi = np.float32(cp[idx][0]), np.float32(cp[nxt][0])
            frac = np.float32(density - lo) / np.float32(hi - lo)
            frac = np.float32(min(max(frac, np.float32(0)), np.float32(1))) if not np.isnan(frac) else np.float32(0)
            v0, v1 = np.array(cp[idx][1:4]), np.array(cp[nxt][1:4])
            out[i] = np.float64(np.float32(1) - frac) * v0 + np.float64(frac) * v1
    return out


def texture_tf_tensor(color_points, plot: Sequence[float], absorption_scaling: float) -> np.ndarray:
    """TransferFunctionTexture::computeTexture: (1,256,4) float32 = colour texture, absorptionScaling * plot[i]."""
    if len(plot) != TEXTURE_RESOLUTION:
        raise RuntimeError(f"a texture TF needs {TEXTURE_RESOLUTION} opacity values, got {len(plot)}")
    out = np.empty((1, TEXTURE_RESOLUTION, 4), np.float32)
    out[0, :, :3] = color_texture(color_points, TEXTURE_RESOLUTION)
    out[0, :, 3] = np.float64(absorption_scaling) * np.asarray(plot, np.float32).astype(np.float64)
    return out


def paste_into_texture(evaluate, absorption_scaling: float = 10.0):
    """TransferFunctionTexture::doPaste of a freshly constructed texture TF (absorptionScaling_ 10, :17): samples `evaluate(d) ->
    rgba` at the 256 texel centres. Returns (colour points, plot, absorption scaling) - feed them to texture_tf_tensor."""
    cps, plot, max_w = [], [], 0.0
    for i in range(TEXTURE_RESOLUTION):
        d = (i + 0.5) / TEXTURE_RESOLUTION
        rgba = [float(v) for v in evaluate(d)]
        cps.append([d, rgba[0], rgba[1], rgba[2]])
        plot.append(float(np.float32(rgba[3] / absorption_scaling)))
        max_w = max(max_w, rgba[3])
    if max_w > absorption_scaling:
        f = absorption_scaling / max_w
        absorption_scaling = max_w
        plot = [float(np.float32(p * f)) for p in plot]
    return cps, plot, absorption_scaling


# ------------------------------------------------------------------------------------------------ the module tree
CHANNELS = {"ChannelMask": 0, "ChannelNormal": 1, "ChannelDepth": 2, "ChannelColor": 3}
ORIENTATIONS = {"Xp": 0, "Xm": 1, "Yp": 2, "Ym": 3, "Zp": 4, "Zm": 5}


def _module(root: dict, tag: str, name: str) -> dict:
    """root[tag][name]; nlohmann's operator[] on a missing key of a const json is undefined - a clear error here."""
    if tag not in root or name not in root[tag]:
        raise RuntimeError(f"settings file has no module '{name}' for tag '{tag}'")
    return root[tag][name]


def _vec3(v, default=(0.0, 0.0, 0.0)):
    if v is None:
        return tuple(default)
    if not isinstance(v, (list, tuple)) or len(v) != 3:
        raise RuntimeError(f"Unable to deserialize {v} into a double3")
    return tuple(float(x) for x in v)


def parse_settings(settings: dict, root_path: str) -> Dict:
    """The loadTree walk as plain data: which modules the file selects and the values their `load` functions read (with the
    defaults those functions use for missing keys). `root_path` resolves a relative volumePath (volume_interpolation_grid.cpp:784-787)."""
    root_name = settings.get("root", "")
    if root_name != "Simple":
        raise RuntimeError(f"image evaluator '{root_name}' is not part of the B200 path (only 'Simple': ImageEvaluatorSimple)")
    ev = _module(settings, "ImageEvaluator", "Simple")
    plan = {"evaluator": {
        # IImageEvaluator::load, iimage_evaluator.cpp:312-317; ImageEvaluatorSimple::load, image_evaluator_simple.cpp:391-411
        "channel": CHANNELS.get(ev.get("outputChannel", ""), 3), "double_precision": bool(ev.get("doublePrecision", False)),
        "samples_per_iteration_log2": int(ev.get("samplesPerIterationLog2", 2)), "use_tonemapping": bool(ev.get("useTonemapping", False)),
        "tonemapping_shoulder": float(ev.get("tonemappingShoulder", 1.0)), "fix_max_exposure": bool(ev.get("fixMaxExposure", False)),
        "fixed_max_exposure": float(ev.get("fixedMaxExposure", 1.0))}}
    if plan["evaluator"]["double_precision"]:
        raise RuntimeError("doublePrecision evaluators are not part of the B200 path")

    cam_name = ev.get("selectedCamera", "")
    if cam_name != "Sphere":
        raise RuntimeError(f"camera '{cam_name}' is not registered (only 'Sphere': CameraOnASphere)")
    cam = _module(settings, "camera", "Sphere")
    orientation = cam.get("orientation", "")
    plan["camera"] = {     # CameraOnASphere::load, camera.cpp:349-361
        "orientation": ORIENTATIONS.get(orientation, ORIENTATIONS["Ym"]), "center": _vec3(cam.get("center")),
        "pitch": float(cam.get("pitch", 0.0)), "yaw": float(cam.get("yaw", 0.0)), "distance": float(cam.get("distance", 0.0)),
        "fov_y": float(cam.get("fovY", math.radians(45.0))), "zoom": float(cam.get("zoom", 1))}

    re_name = ev.get("selectedRayEvaluator", "")
    if re_name != "DVR":
        raise RuntimeError(f"ray evaluator '{re_name}' is not part of the B200 path (only 'DVR': RayEvaluationSteppingDvr)")
    dvr = _module(settings, "RayEvaluation", "DVR")
    stepsize = float(dvr.get("stepsize", 0.005))
    if dvr.get("stepsizeIsObjectSpace", False):
        stepsize /= 256         # "fix for old config files that stored the stepsize in object space", ray_evaluation_stepping.cpp:66-71
    plan["dvr"] = {"stepsize": stepsize, "min_density": float(dvr.get("minDensity", 0.0)), "max_density": float(dvr.get("maxDensity", 1.0)),
                   "early_out": bool(dvr.get("earlyOut", True))}       # RayEvaluationSteppingDvr::load, :740-755

    # volume: at the evaluator in newer files, at the ray evaluator in older ones (IRayEvaluation::load, ray_evaluation.cpp:17-33;
    # loaded after the evaluator, so it wins when both name one)
    vol_name = dvr.get("selectedVolume", "") or ev.get("selectedVolume", "")
    if vol_name == "Grid":
        g = _module(settings, "volume", "Grid")
        source = g.get("source", "")
        if source not in ("VOLUME", "EMPTY", ""):
            raise RuntimeError(f"volume source '{source}' is not part of the B200 path (only VOLUME: a .cvol file)")
        interp = g.get("interpolation", "") or "TRILINEAR"
        if interp not in ("TRILINEAR", "NEAREST_NEIGHBOR", "TRICUBIC"):
            interp = "TRILINEAR"
        path = None
        if source == "VOLUME":
            path = g.get("volumePath", "")
            if not os.path.isabs(path):
                path = os.path.abspath(os.path.join(root_path, path))
            path = os.path.normpath(path)
        plan["volume"] = {"kind": "Grid", "path": path, "interpolation": interp, "mipmap_level": int(g.get("mipmapLevel", 0)),
                          "density_feature": int(g.get("selectedDensityFeatureIndex", 0))}
    elif vol_name == "SRN":                       # VolumeInterpolationNetwork::getName, volume_interpolation_network.cpp:1522-1525
        n = settings.get("volume", {}).get(vol_name, {})       # ::load, :1664-1672 (the networks themselves are not stored: "TODO" there)
        gm = {"OFF_OR_DIRECT": 0, "FINITE_DIFFERENCES": 1, "ADJOINT_METHOD": 2}.get(n.get("gradientMode", ""), 0)
        plan["volume"] = {"kind": "Network", "gradient_mode": gm, "finite_differences_stepsize": float(n.get("finiteDifferencesStepsize", 1 / 256.0))}
    elif vol_name == "":
        plan["volume"] = None
    else:
        raise RuntimeError(f"volume module '{vol_name}' is not part of the B200 path")

    tf_name = dvr.get("selectedTF", "")
    if tf_name == "Piecewise":
        t = _module(settings, "tf", "Piecewise")
        plan["tf"] = {"kind": "Piecewise", "color_points": t["colorPoints"], "opacity_points": t["opacityPoints"],
                      "absorption_scaling": float(t["absorptionScaling"])}
    elif tf_name == "Texture":
        t = _module(settings, "tf", "Texture")
        pre = t.get("preintegrationMode", "") or "None"
        plan["tf"] = {"kind": "Texture", "color_points": t["colorPoints"], "plot": [float(v) for v in t["opacityPoints"]],
                      "absorption_scaling": float(t["absorptionScaling"]), "preintegration_mode": {"None": 0, "Preintegrate1D": 1, "Preintegrate2D": 2}.get(pre, 0)}
    elif tf_name == "Identity":
        t = _module(settings, "tf", "Identity")
        plan["tf"] = {"kind": "Identity", "absorption_scaling": float(t.get("absorptionScaling", 1.0)), "emission_scaling": float(t.get("emissionScaling", 1.0))}
    elif tf_name == "":
        plan["tf"] = None
    else:
        raise RuntimeError(f"transfer function '{tf_name}' is not part of the B200 path (Piecewise, Texture, Identity are)")

    brdf_name = dvr.get("selectedBRDF", "")
    if brdf_name == "Lambert":
        b = _module(settings, "brdf", "Lambert")        # BRDFLambert::load, brdf.cpp:208-225
        plan["brdf"] = {"enable_magnitude_scaling": bool(b.get("enableMagnitudeScaling", False)), "magnitude_scaling": float(b.get("magnitudeScaling", 1.0)),
                        "enable_phong": bool(b.get("enablePhong", False)), "ambient": float(b.get("ambient", 1.0)), "specular": float(b.get("specular", 1.0)),
                        "magnitude_center": float(b.get("magnitudeCenter", 1.0)), "magnitude_radius": float(b.get("magnitudeRadius", 1.0)),
                        "specular_exponent": int(b.get("specularExponent", 1)), "light_follows_camera": bool(b.get("lightFollowsCamera", True)),
                        "light_type": {"Point": 0, "Directional": 1}.get(b.get("lightType", ""), 1),
                        "light_position": _vec3(b.get("lightPosition")), "light_direction": _vec3(b.get("lightDirection"))}
    elif brdf_name == "":
        plan["brdf"] = None
    else:
        raise RuntimeError(f"BRDF '{brdf_name}' is not registered (only 'Lambert')")
    # The blending module is the registry's single instance, not part of the tree loadTree walks (ray_evaluation_stepping.cpp:540-544):
    # load_from_json leaves it at its default (BeerLambert) whatever the file says. Reported so that a caller can apply it.
    plan["file_blending"] = settings.get("blending", {}).get("blending", {}).get("blending", None)
    return plan


def read_settings(filename: str) -> Dict:
    if not os.path.isfile(filename):
        raise RuntimeError("Unable to open file")
    with open(filename, "r") as f:
        try:
            settings = json.load(f)
        except json.JSONDecodeError as e:
            raise RuntimeError(f"Unable to load json with the settings ({filename}): {e}")
    return parse_settings(settings, os.path.dirname(os.path.abspath(filename)))
