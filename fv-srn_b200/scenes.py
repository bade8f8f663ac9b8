"""
Synthetic scenes for parity tests and the benchmark: random-init networks of the reference's
architecture, the camera-on-a-sphere reference frame, and the piecewise transfer function of
applications/config-files/plume100-v1-dvr.json. NumPy only.
"""
from __future__ import annotations

import math

import numpy as np

from . import volnet

# ------------------------------------------------------------------------------------------------ camera
_UP = [(1, 0, 0), (-1, 0, 0), (0, 1, 0), (0, -1, 0), (0, 0, 1), (0, 0, -1)]
_PERM = [(2, -1, -3), (-2, 1, 3), (1, 2, 3), (-1, -2, -3), (-3, -1, 2), (3, 1, -2)]
_INV_YAW = [False, True, True, False, True, False]
ORIENTATIONS = ["Xp", "Xm", "Yp", "Ym", "Zp", "Zm"]


def camera_reference_frame(pitch, yaw, distance, center=(0.0, 0.0, 0.0), orientation="Ym") -> np.ndarray:
    """CameraOnASphere -> (1,3,3) float32 rows eye/right/up.
    eulerToCartesian + computeParameters + updateCameraMatrix, reference renderer/camera.cpp:480-509, 553-581."""
    o = ORIENTATIONS.index(orientation) if isinstance(orientation, str) else int(orientation)
    yaw = yaw if _INV_YAW[o] else -yaw
    pitch = -pitch  # OrientationInvertPitch is false for every orientation (camera.cpp:33-35)
    pos = (math.cos(pitch) * math.cos(yaw) * distance, math.sin(pitch) * distance, math.cos(pitch) * math.sin(yaw) * distance)
    origin = np.array([pos[abs(p) - 1] * (1 if p > 0 else -1) for p in _PERM[o]], np.float64) + np.asarray(center, np.float64)
    up = np.asarray(_UP[o], np.float64)
    front = np.asarray(center, np.float64) - origin
    front /= np.linalg.norm(front)
    right = np.cross(front, up); right /= np.linalg.norm(right)
    up2 = np.cross(right, front); up2 /= np.linalg.norm(up2)
    return np.stack([origin, right, up2]).astype(np.float32)[None]


# values named in SURVEY §8d (plume100-v1-dvr.json pitch/yaw/fov, ejecta70-v6-dvr.json distance)
DEFAULT_CAMERA = dict(pitch=0.2618, yaw=2.3038, distance=1.6105, center=(0.0, 0.0, 0.0), orientation="Ym")
DEFAULT_FOV_Y = 0.7853981633974483


def camera_sweep(n_views: int, **kw) -> np.ndarray:
    """n_views cameras on a full rotation (inference.py:396) -> (n,3,3)."""
    base = dict(DEFAULT_CAMERA); base.update(kw)
    yaw0 = base.pop("yaw")
    return np.concatenate([camera_reference_frame(yaw=yaw0 + a, **base) for a in np.linspace(0, 2 * np.pi, n_views, endpoint=False)])


def estimate_row_cost(camera: np.ndarray, fov_y: float, width: int, height: int, box_min=(-0.5, -0.5, -0.5),
                      box_size=(1.0, 1.0, 1.0), sub: int = 16) -> np.ndarray:
    """Per-row march cost of one view for screen-band sharding (dist.partition_rows_by_cost): the box-intersection length
    tmax - max(tmin, 0) of every `sub`-th ray (CameraReferenceFrame::eval + intersectionRayAABB, renderer_camera.cuh:41-51,
    renderer_utils.cuh:91-105, in NumPy), summed over x and spread over the rows it stands for. -> (height,) float64."""
    cam = np.asarray(camera, np.float64).reshape(-1, 3, 3)[0]
    eye, right, up = cam
    front = np.cross(up, right)
    tan_y = math.tan(fov_y / 2); tan_x = tan_y * (width / height)
    xs = np.arange(sub // 2, width, sub); ys = np.arange(sub // 2, height, sub)
    nx = 2 * (xs + 0.5) / width - 1; ny = 2 * (ys + 0.5) / height - 1
    d = front[None, None] + (nx * tan_x)[None, :, None] * right[None, None] + (ny * tan_y)[:, None, None] * up[None, None]
    d /= np.linalg.norm(d, axis=-1, keepdims=True)
    with np.errstate(divide="ignore", invalid="ignore"):
        inv = 1.0 / d
        t0 = (np.asarray(box_min) - eye) * inv; t1 = (np.asarray(box_min) + np.asarray(box_size) - eye) * inv
    tmin = np.maximum(np.minimum(t0, t1).max(-1), 0.0); tmax = np.maximum(t0, t1).min(-1)
    length = np.maximum(tmax - tmin, 0.0).sum(1)                  # per sampled row
    cost = np.zeros(height)
    for i, y in enumerate(ys):
        y0 = y - sub // 2
        cost[y0:min(y0 + sub, height)] = length[i] / sub
    return cost


# ------------------------------------------------------------------------------------------------ transfer function
PLUME_COLOR_POINTS = [
    (0.0033444815780967474, (9.999999974752427e-07, 9.999899930335232e-07, 9.999899930335232e-07)),
    (0.44147157669067383, (1.0, 0.039215683937072754, 0.039215683937072754)),
    (0.7558528184890747, (0.9166666865348816, 0.9166665077209473, 0.11683008819818497)),
    (0.9832776188850403, (1.0, 1.0, 0.9999899864196777)),
]
PLUME_OPACITY_POINTS = [(0.0568561889231205, 0.0), (0.3645485043525696, 0.6599999666213989), (1.0, 0.8999999761581421)]
PLUME_ABSORPTION_SCALING = 25.37


def piecewise_tf_tensor(color_points=PLUME_COLOR_POINTS, opacity_points=PLUME_OPACITY_POINTS,
                        absorption_scaling=PLUME_ABSORPTION_SCALING) -> np.ndarray:
    """TransferFunctionPiecewiseLinear::computeTensor (transfer_function_piecewise.cpp:166-281) -> (1,R,5) float32."""
    cp = [(float(p), np.asarray(c, np.float64)) for p, c in color_points]
    op = [(float(p), float(a)) for p, a in opacity_points]
    if cp[0][0] > 0: cp.insert(0, (-1.0, cp[0][1]))
    if op[0][0] > 0: op.insert(0, (-1.0, op[0][1]))
    if cp[-1][0] < 1: cp.append((2.0, cp[-1][1]))
    if op[-1][0] < 1: op.append((2.0, op[-1][1]))
    pts = [[min(cp[0][0], op[0][0]) if cp[0][0] <= op[0][0] else op[0][0], cp[0][1], op[0][1]]]
    io = ic = 0
    while io < len(op) - 1 and ic < len(cp) - 1:
        if op[io + 1][0] < cp[ic + 1][0]:
            f = (op[io + 1][0] - cp[ic][0]) / (cp[ic + 1][0] - cp[ic][0])
            pts.append([op[io + 1][0], cp[ic][1] + f * (cp[ic + 1][1] - cp[ic][1]), op[io + 1][1]])
            io += 1
        else:
            f = (cp[ic + 1][0] - op[io][0]) / (op[io + 1][0] - op[io][0])
            pts.append([cp[ic + 1][0], cp[ic + 1][1], op[io][1] + f * (op[io + 1][1] - op[io][1])])
            ic += 1
    eps = 1e-7
    i = 0
    while i < len(pts) - 2:
        if (pts[i][2] < eps and pts[i + 1][2] < eps and pts[i + 2][2] < eps) or (pts[i + 1][0] - pts[i][0] < eps):
            del pts[i + 1]
        else:
            i += 1
    out = np.zeros((1, len(pts), 5), np.float32)
    for i, (pos, rgb, a) in enumerate(pts):
        out[0, i, :3] = np.clip(rgb, 0.0, 1.0 - np.finfo(np.float32).eps)
        out[0, i, 3] = min(max(a, 0.0), 1.0) * absorption_scaling
        out[0, i, 4] = pos
    return out


def texture_tf_from_piecewise(tf: np.ndarray, resolution: int = 256) -> np.ndarray:
    """Samples a piecewise TF at texel centres -> (1,R,4) (what convert_to_texture_tf does conceptually)."""
    R = tf.shape[1]
    out = np.zeros((1, resolution, 4), np.float32)
    for k in range(resolution):
        d = (k + 0.5) / resolution
        i = 0
        while i < R - 2 and not tf[0, i + 1, 4] > d:
            i += 1
        p0, p1 = tf[0, i, 4], tf[0, i + 1, 4]
        f = (min(max(d, p0), p1) - p0) / (p1 - p0)
        out[0, k] = tf[0, i, :4] + f * (tf[0, i + 1, :4] - tf[0, i, :4])
    return out


# ------------------------------------------------------------------------------------------------ networks
def make_network(hidden: int = 32, num_layers: int = 4, grid_res: int = 32, grid_channels: int = 16, activation: str = "ReLU",
                 fourier: str = "nerf", fourier_std: float = 1.0, output_mode: str = "density:direct", seed: int = 42,
                 grid_std: float = 0.5, time_keyframes: int = 0, has_direction: bool = False, pass_time: bool = False,
                 box_min=(-0.5, -0.5, -0.5), box_size=(1.0, 1.0, 1.0), last_gain: float = 1.0,
                 last_bias: float | None = None, hidden_gain: float = 1.0, direction_in_fourier: bool = False) -> volnet.SceneNetwork:
    """Random-init SRN shaped like the reference's SceneRepresentationNetwork (network.py:653-770):
    ``num_layers`` linear layers (``--layers h:h:..`` has num_layers-1 entries), Fourier count (hidden-4)/2
    (collect_possible_layers.py:50), nn.Linear default init U(-1/sqrt(fan_in), 1/sqrt(fan_in)), grid ~ N(0, grid_std).
    time_keyframes > 0 creates that many time-dependent latent grids (volumetric_features_time_dependent).
    last_gain / last_bias scale the last layer's weights and replace its bias: a random-init network's output is a narrow
    band around some mean (deep ReLU networks more so), these two place that band across the transfer function so
    that the synthetic frame has content (the same reason the latent grid uses std 0.5 instead of 0.01).
    hidden_gain scales the weights of every layer but the last: nn.Linear's default init shrinks the signal by ~0.4 per
    ReLU layer, so a 6-layer random network's output varies by 1e-2 of its activations and only a large last_gain (which
    amplifies fp16 noise as much as signal) would make it visible; hidden_gain = sqrt(6) (He-uniform) keeps the signal
    O(1) through the layers, like trained weights do."""
    rng = np.random.default_rng(seed)
    foff = 8 if has_direction else 4
    nf = (hidden - foff) // 2
    net = volnet.SceneNetwork()
    net.has_direction = has_direction
    net.has_time = bool(pass_time)
    fc = 6 if (has_direction and direction_in_fourier) else 3     # useDirectionInFourierFeatures: B multiplies (position, direction)
    if fourier == "nerf":  # scaled block identity (network.py:55-63)
        blocks = [(2.0 ** i) * np.eye(fc) for i in range(int(math.ceil(nf / fc)))]
        B = np.concatenate(blocks, 0)[:nf]
    else:
        B = rng.normal(0.0, fourier_std, (nf, fc))
    net.set_fourier_matrix(B * (2 * np.pi), premultiplied=True)
    net.output_mode = output_mode
    cout = volnet.OUTPUT_CHANNELS_IN[volnet.OUTPUT_MODES.index(output_mode)]
    act_name, _, act_param = activation.partition(":")
    act_param = float(act_param) if act_param else 1.0
    # torch-side input order: [pos(3), (time), (dir 3), cos, sin, latent]
    cin = 3 + (1 if pass_time else 0) + (3 if has_direction else 0) + 2 * nf + grid_channels
    sizes = [hidden] * (num_layers - 1)
    last = cin
    for s in sizes:
        k = 1.0 / math.sqrt(last)
        net.add_layer(rng.uniform(-k, k, (s, last)) * hidden_gain, rng.uniform(-k, k, (s,)), act_name, act_param)
        last = s
    k = 1.0 / math.sqrt(last)
    w_last = rng.uniform(-k, k, (cout, last)) * last_gain
    b_last = rng.uniform(-k, k, (cout,))
    if last_bias is not None:
        b_last = np.full((cout,), float(last_bias))
    if output_mode.startswith("rgbo"):
        b_last = np.abs(b_last) + 1.0  # network.py:405-406
    net.add_layer(w_last, b_last, "None", 1.0)
    net.box_min, net.box_size = tuple(box_min), tuple(box_size)
    if grid_channels > 0:
        net.has_latent_grid = True
        T = max(1, time_keyframes)
        net.time_min, net.time_num, net.time_step, net.ensemble_min, net.ensemble_num = 0, T, 1, 0, 0
        for _ in range(T):
            g = (rng.standard_normal((grid_channels, grid_res, grid_res, grid_res)) * grid_std).astype(np.float32)
            net.time_grids.append(volnet.LatentGrid(volnet.GRID_FLOAT, g))
    assert net.valid(), "generated network is not valid"
    return net


def encode_grids(net: volnet.SceneNetwork, encoding: int) -> volnet.SceneNetwork:
    """The same network with its latent grids byte-encoded (LatentGrid::Encoding 1 = linear, 2 = gaussian) by the library's
    own encoder (fvsrn_network_set_*_grid = LatentGrid::initEncodingByte*, volume_interpolation_network.cpp:336-431)."""
    import ctypes as C
    from . import _lib
    L = _lib.lib()
    data = volnet.dumps(net)
    h = C.c_void_p()
    _lib.check(L.fvsrn_network_load_volnet(data, len(data), C.byref(h)))
    try:
        for i, g in enumerate(net.time_grids):
            a = np.ascontiguousarray(g.values, np.float32)
            _lib.check(L.fvsrn_network_set_time_grid(h, i, a.ctypes.data, *a.shape, int(encoding), None))
        for i, g in enumerate(net.ensemble_grids):
            a = np.ascontiguousarray(g.values, np.float32)
            _lib.check(L.fvsrn_network_set_ensemble_grid(h, i, a.ctypes.data, *a.shape, int(encoding), None))
        n = C.c_size_t()
        _lib.check(L.fvsrn_network_save_volnet(h, None, 0, C.byref(n)))
        buf = C.create_string_buffer(n.value)
        _lib.check(L.fvsrn_network_save_volnet(h, buf, n.value, C.byref(n)))
        return volnet.loads(buf.raw[:n.value])
    finally:
        L.fvsrn_network_free(h)


# BASELINE.json configs (C1..C5)
CONFIGS = {
    "C1": dict(net=dict(hidden=32, num_layers=4, grid_res=16), width=256, height=256, stepsize=0.005, views=1, spp=1),
    "C2": dict(net=dict(hidden=32, num_layers=4, grid_res=32), positions=2 ** 24),
    "C3": dict(net=dict(hidden=32, num_layers=4, grid_res=32), width=1024, height=1024, stepsize=1.0 / 512, views=1, spp=1),
    "C4": dict(net=dict(hidden=64, num_layers=6, grid_res=64), width=2048, height=2048, stepsize=1.0 / 512, views=1, spp=4),
    "C5": dict(net=dict(hidden=32, num_layers=4, grid_res=32, time_keyframes=8), width=512, height=512, stepsize=1.0 / 512, views=64, spp=1),
}
