"""
ctypes front end of the CPU oracle (oracle/fvsrn_oracle.c) — TEST INFRASTRUCTURE ONLY.

May be imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs,
never by the product package. See the header of fvsrn_oracle.c for what it restates and how it is pinned.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from math import floor

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "_build", "libfvsrn_oracle.so")
_lib = None

ACTIVATIONS = ["ReLU", "Sine", "Snake", "SnakeAlt", "Sigmoid", "None"]
OUTPUT_MODES = ["density", "density:direct", "rgbo", "rgbo:direct", "densitygrad", "densitygrad:direct", "densitygrad:cubic",
                "densitycurvature", "densitycurvature:direct"]


def build(force: bool = False):
    src = os.path.join(_HERE, "fvsrn_oracle.c")
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE], stdout=subprocess.DEVNULL)
    return _LIB_PATH


class _Net(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("H", "nf", "has_dir", "dir_in_fourier", "has_time", "grid_c", "num_hidden",
                                          "activation", "output_mode", "cout")] + [
        ("act_param", C.c_float), ("time_value", C.c_float),
        ("fourier", C.c_void_p), ("w0", C.c_void_p), ("b0", C.c_void_p), ("wh", C.c_void_p), ("bh", C.c_void_p),
        ("wlast", C.c_void_p), ("blast", C.c_void_p),
        ("ngroups", C.c_int32), ("gx", C.c_int32), ("gy", C.c_int32), ("gz", C.c_int32),
        ("ga", C.c_void_p * 4), ("gb", C.c_void_p * 4), ("frac", C.c_float * 4), ("gstride", C.c_int32 * 4),
        ("box_min", C.c_float * 3), ("box_size", C.c_float * 3), ("tex_filter", C.c_int32), ("acc32", C.c_int32),
        ("encoding", C.c_int32), ("b_uses_a", C.c_int32),
        ("enc_off_a", C.c_float * 64), ("enc_scale_a", C.c_float * 64), ("enc_off_b", C.c_float * 64), ("enc_scale_b", C.c_float * 64)]


class _Dvr(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("width", "height", "batch", "tile_x0", "tile_y0", "tile_w", "tile_h")] + [
        ("camera", C.c_void_p), ("fov_y", C.c_float), ("stepsize", C.c_float), ("alpha_early_out", C.c_float),
        ("density_min", C.c_float), ("density_max", C.c_float), ("early_out", C.c_int32), ("blend_mode", C.c_int32),
        ("tf_kind", C.c_int32), ("tf_R", C.c_int32), ("tf_batch", C.c_int32), ("tf", C.c_void_p), ("spp", C.c_int32),
        ("fp32_mode", C.c_int32),
        ("gradient_mode", C.c_int32), ("fd_stepsize", C.c_float), ("brdf_flags", C.c_int32), ("light_type", C.c_int32),
        ("light", C.c_float * 3), ("brdf_magnitude_scaling", C.c_float), ("brdf_ambient", C.c_float), ("brdf_specular", C.c_float),
        ("brdf_magnitude_center", C.c_float), ("brdf_magnitude_radius", C.c_float), ("brdf_specular_exponent", C.c_int32)]


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_LIB_PATH)
        _lib.oracle_evaluate.argtypes = [C.POINTER(_Net), C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int]
        _lib.oracle_features.argtypes = [C.POINTER(_Net), C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64]
        _lib.oracle_render_dvr.argtypes = [C.POINTER(_Net), C.POINTER(_Dvr), C.c_void_p, C.c_void_p, C.c_void_p]
    return _lib


class OracleNetwork:
    """Flattens a ``volnet.SceneNetwork`` (parsed .volnet) into the oracle's C struct."""

    def __init__(self, net, time: float = 0.0, ensemble: int = 0, unit_box: bool = False, tex_filter: bool = False, acc32: bool = False,
                 b_uses_a: bool = False):
        self.net = net
        self._keep = []
        n = _Net()
        n.acc32 = int(acc32)             # fp32 accumulation / activation per layer (tcgen05 kernels) instead of fp16 (wmma path)
        n.tex_filter = int(tex_filter)   # latent fetch with texture-unit weight quantisation (reference / tcgen05 kernel)
        H = net.input_channels_out()
        gc = net.latent_channels()
        assert net.num_fourier > 0, "oracle: networks without Fourier features are not on the hot path"
        L = len(net.layers)
        n.H, n.nf, n.has_dir, n.dir_in_fourier = H, net.num_fourier, int(net.has_direction), int(net.direction_in_fourier)
        n.has_time, n.grid_c, n.num_hidden = int(net.has_time), gc, L - 2
        n.activation = ACTIVATIONS.index(net.layers[0].activation)
        n.act_param = float(net.layers[0].activation_param)
        n.output_mode = OUTPUT_MODES.index(net.output_mode)
        n.cout = net.layers[-1].cout

        def keep(a, dtype=np.float16):
            a = np.ascontiguousarray(a, dtype=dtype); self._keep.append(a)
            return a.ctypes.data

        n.fourier = keep(net.fourier.T)                       # [cin][nf]
        n.w0 = keep(net.layers[0].weights); n.b0 = keep(net.layers[0].bias)
        if L > 2:
            n.wh = keep(np.stack([l.weights for l in net.layers[1:-1]])); n.bh = keep(np.stack([l.bias for l in net.layers[1:-1]]))
        n.wlast = keep(net.layers[-1].weights); n.blast = keep(net.layers[-1].bias)
        n.time_value = 0.0
        g = 0
        if net.has_latent_grid:
            enc = (net.time_grids or net.ensemble_grids)[0].encoding
            n.encoding, n.b_uses_a = enc, int(b_uses_a)
            bpe = 1 if enc else 2       # bytes per stored element

            def chlast(grid):  # (C,Z,Y,X) -> (Z,Y,X,C): fp16 values, or the stored bytes of a byte-encoded grid
                src = grid.raw.astype(np.uint8) if enc else grid.values.astype(np.float32).astype(np.float16)
                a = np.ascontiguousarray(np.transpose(src, (1, 2, 3, 0)))
                self._keep.append(a); return a

            def consts(gidx, A, B, i):
                if not enc: return
                for c in range(16):
                    n.enc_off_a[16 * gidx + c] = A.offset_or_mean[16 * i + c]; n.enc_scale_a[16 * gidx + c] = A.scale_or_std[16 * i + c]
                    n.enc_off_b[16 * gidx + c] = B.offset_or_mean[16 * i + c]; n.enc_scale_b[16 * gidx + c] = B.scale_or_std[16 * i + c]
            if net.time_grids:
                t = float(np.float32(min(max((np.float32(time) - np.float32(net.time_min)) / np.float32(net.time_step), 0.0), float(net.time_num - 1))))
                lo = min(int(t), net.time_num - 1); hi = min(lo + 1, net.time_num - 1)
                frac = 0.0 if lo == hi else float(np.float32(t) - np.float32(floor(t)))
                n.time_value = t
                A, B = chlast(net.time_grids[lo]), chlast(net.time_grids[hi])
                Cg = A.shape[3]; n.gz, n.gy, n.gx = A.shape[0], A.shape[1], A.shape[2]
                for i in range(Cg // 16):
                    n.ga[g] = A.ctypes.data + 16 * bpe * i; n.gb[g] = B.ctypes.data + 16 * bpe * i
                    consts(g, net.time_grids[lo], net.time_grids[hi], i)
                    n.frac[g] = frac; n.gstride[g] = Cg; g += 1
            if net.ensemble_grids:
                e = min(max(ensemble - net.ensemble_min, 0), net.ensemble_num - 1)
                A = chlast(net.ensemble_grids[e])
                Cg = A.shape[3]; n.gz, n.gy, n.gx = A.shape[0], A.shape[1], A.shape[2]
                for i in range(Cg // 16):
                    n.ga[g] = A.ctypes.data + 16 * bpe * i; n.gb[g] = A.ctypes.data + 16 * bpe * i
                    consts(g, net.ensemble_grids[e], net.ensemble_grids[e], i)
                    n.frac[g] = 0.0; n.gstride[g] = Cg; g += 1
        n.ngroups = g
        for i in range(3):
            n.box_min[i] = 0.0 if unit_box else net.box_min[i]
            n.box_size[i] = 1.0 if unit_box else net.box_size[i]
        self.c = n

    @property
    def cout(self): return int(self.c.cout)


def evaluate(net, positions, directions=None, time=0.0, ensemble=0, fp32=False, tex_filter=False, acc32=False):
    """IVolumeInterpolation::evaluate restated on the CPU. positions (N,3) in [0,1]^3 -> (N,cout) float32.
    tex_filter: fetch the latent grid with the texture unit's 8-bit fractional weights (reference + tcgen05 kernel)."""
    on = OracleNetwork(net, time, ensemble, unit_box=True, tex_filter=tex_filter, acc32=acc32)
    pos = np.ascontiguousarray(positions, dtype=np.float32)
    d = None if directions is None else np.ascontiguousarray(directions, dtype=np.float32)
    out = np.zeros((pos.shape[0], on.cout), np.float32)
    lib().oracle_evaluate(C.byref(on.c), pos.ctypes.data, None if d is None else d.ctypes.data, out.ctypes.data,
                          pos.shape[0], int(fp32))
    # densitygrad* / densitycurvature* networks: `evaluate` is the density channel (OutputModeNumChannelsOut = 1)
    return out[:, :1].copy() if on.c.output_mode >= 4 else out


def adjoint_gradient(net, positions, time=0.0, ensemble=0, tex_filter=False, latent_step=None):
    """_evalNormalAdjoint restated on the CPU (fvsrn_oracle.c adjoint_h): positions (N,3) in the unit box -> d density / d position (N,3)
    in unit-box coordinates. latent_step: central-difference step through the latent grid, default 1 / (4 * resolution)."""
    on = OracleNetwork(net, time, ensemble, unit_box=True, tex_filter=tex_filter, acc32=False)
    pos = np.ascontiguousarray(positions, dtype=np.float32)
    h = float(np.float32(1.0) / (np.float32(on.c.gx) * np.float32(4.0))) if latent_step is None else float(latent_step)
    out = np.zeros((pos.shape[0], 3), np.float32)
    lib().oracle_adjoint_gradient.argtypes = [C.c_void_p, C.c_void_p, C.c_float, C.c_void_p, C.c_int64]
    lib().oracle_adjoint_gradient(C.byref(on.c), pos.ctypes.data, h, out.ctypes.data, pos.shape[0])
    return out


def features(net, positions, time=0.0, ensemble=0, tex_filter=False, acc32=False):
    on = OracleNetwork(net, time, ensemble, unit_box=True, tex_filter=tex_filter, acc32=acc32)
    pos = np.ascontiguousarray(positions, dtype=np.float32)
    K0 = on.c.H + on.c.grid_c
    rows = np.empty((pos.shape[0], K0), np.float16)
    lib().oracle_features(C.byref(on.c), pos.ctypes.data, None, rows.ctypes.data, pos.shape[0])
    return rows


def render_dvr(net, *, width, height, camera, fov_y, stepsize, tf, tf_kind=0, alpha_early_out=1 - 1e-5,
               density_min=0.0, density_max=1.0, early_out=True, blend_mode=1, spp=1, time=0.0, ensemble=0,
               tile=None, fp32=False, tex_filter=False, acc32=False, fd_stepsize=None, brdf=None, adjoint=False):
    """ImageEvaluatorSimple::render restated on the CPU -> (img (B,8,H,W) f32, steps (B,H,W) i32, valid_samples).
    fd_stepsize: finite-difference normals (GRADIENT_MODE 1); brdf: dict(flags=("magnitude","phong"), light_type, light,
    magnitude_scaling, ambient, specular, magnitude_center, magnitude_radius, specular_exponent) = BRDFLambert."""
    on = OracleNetwork(net, time, ensemble, unit_box=False, tex_filter=tex_filter, acc32=acc32)
    cam = np.ascontiguousarray(camera, dtype=np.float32).reshape(-1, 3, 3)
    B = cam.shape[0]
    d = _Dvr()
    d.width, d.height, d.batch = width, height, B
    if tile is not None:
        d.tile_x0, d.tile_y0, d.tile_w, d.tile_h = tile
    d.camera = cam.ctypes.data; d.fov_y = fov_y
    d.stepsize, d.alpha_early_out, d.density_min, d.density_max = stepsize, alpha_early_out, density_min, density_max
    d.early_out, d.blend_mode, d.spp, d.fp32_mode = int(early_out), blend_mode, spp, int(fp32)
    if fd_stepsize is not None:
        d.gradient_mode, d.fd_stepsize = 1, fd_stepsize
    if adjoint:     # GRADIENT_MODE 2; the step through the latent grid is 1 / (4 * resolution) (volume_interpolation_network.cpp:1424, 1809-1812)
        d.gradient_mode, d.fd_stepsize = 2, float(np.float32(1.0) / (np.float32(on.c.gx) * np.float32(4.0)))
    if brdf is not None:
        d.brdf_flags = (1 if "magnitude" in brdf["flags"] else 0) | (2 if "phong" in brdf["flags"] else 0)
        d.light_type = brdf["light_type"]; d.light = (C.c_float * 3)(*brdf["light"])
        d.brdf_magnitude_scaling, d.brdf_ambient, d.brdf_specular = brdf["magnitude_scaling"], brdf["ambient"], brdf["specular"]
        d.brdf_magnitude_center, d.brdf_magnitude_radius = brdf["magnitude_center"], brdf["magnitude_radius"]
        d.brdf_specular_exponent = brdf["specular_exponent"]
    tfa = None
    if tf is not None:
        tfa = np.ascontiguousarray(tf, dtype=np.float32)
        if tf_kind == 2:
            d.tf_R, d.tf_batch = 0, 1
        else:
            tfa = tfa.reshape(-1, tfa.shape[-2], tfa.shape[-1])
            d.tf_R, d.tf_batch = tfa.shape[1], tfa.shape[0]
        d.tf = tfa.ctypes.data
    d.tf_kind = tf_kind
    img = np.zeros((B, 8, height, width), np.float32)
    steps = np.zeros((B, height, width), np.int32)
    counters = np.zeros(2, np.uint64)
    lib().oracle_render_dvr(C.byref(on.c), C.byref(d), img.ctypes.data, steps.ctypes.data, counters.ctypes.data)
    return img, steps, int(counters[0])
