"""
Pure-Python/NumPy reader and writer of the ``.volnet`` binary format.

Follows the reference loaders/savers in renderer/volume_interpolation_network.cpp:
InputParametrization :68-127, OutputParametrization :205-221, Layer :250-288, LatentGrid :564-614,
LatentGridTimeAndEnsemble :756-796, SceneNetwork :1057-1104, and the layer re-layout of
SceneNetwork::addLayer :806-894.

This module is host-side glue (fixtures, tests, building networks from torch modules); the product path
parses ``.volnet`` natively in C++ (csrc/network.cpp) behind ``fvsrn_network_load_volnet``.
"""
from __future__ import annotations

import struct
from dataclasses import dataclass, field
from typing import List, Optional

import numpy as np

ACTIVATIONS = ["ReLU", "Sine", "Snake", "SnakeAlt", "Sigmoid", "None"]
OUTPUT_MODES = ["density", "density:direct", "rgbo", "rgbo:direct", "densitygrad", "densitygrad:direct",
                "densitygrad:cubic", "densitycurvature", "densitycurvature:direct"]
OUTPUT_CHANNELS_IN = [1, 1, 4, 4, 4, 4, 4, 6, 6]
OUTPUT_CHANNELS_OUT = [1, 1, 4, 4, 1, 1, 1, 1, 1]
GRID_FLOAT, GRID_BYTE_LINEAR, GRID_BYTE_GAUSSIAN = 0, 1, 2


@dataclass
class Layer:
    """One linear layer; ``weights`` is canonical (cout, cin) float16 in the *kernel's* input order."""
    weights: np.ndarray
    bias: np.ndarray
    activation: str = "None"
    activation_param: float = 1.0

    @property
    def cout(self): return self.weights.shape[0]

    @property
    def cin(self): return self.weights.shape[1]


@dataclass
class LatentGrid:
    """Dense (C,Z,Y,X) float32 values (decoded) plus, for byte encodings, the stored bytes."""
    encoding: int
    values: np.ndarray                       # (C,Z,Y,X) float32, decoded
    raw: Optional[np.ndarray] = None         # (C,Z,Y,X) uint8 for byte encodings
    offset_or_mean: Optional[np.ndarray] = None
    scale_or_std: Optional[np.ndarray] = None

    @property
    def shape(self): return self.values.shape


@dataclass
class SceneNetwork:
    has_time: bool = False
    has_direction: bool = False
    direction_in_fourier: bool = False
    fourier: Optional[np.ndarray] = None     # (nf, cin) float16, already multiplied by 2*pi
    output_mode: str = "density"
    layers: List[Layer] = field(default_factory=list)
    box_min: tuple = (-5.0, -5.0, -5.0)
    box_size: tuple = (1.0, 1.0, 1.0)
    has_latent_grid: bool = False
    time_min: int = 0
    time_num: int = 0
    time_step: int = 1
    ensemble_min: int = 0
    ensemble_num: int = 0
    time_grids: List[LatentGrid] = field(default_factory=list)
    ensemble_grids: List[LatentGrid] = field(default_factory=list)

    @property
    def num_fourier(self): return 0 if self.fourier is None else int(self.fourier.shape[0])

    def input_channels_out(self):
        if self.num_fourier > 0:
            return 4 + (4 if self.has_direction else 0) + 2 * self.num_fourier
        return 3 + (3 if self.has_direction else 0)

    def latent_channels(self):
        if not self.has_latent_grid:
            return 0
        c = self.time_grids[0].shape[0] if self.time_grids else 0
        c += self.ensemble_grids[0].shape[0] if self.ensemble_grids else 0
        return c

    # -- builder mirroring pyrenderer.SceneNetwork.add_layer (addLayerFromTorch :896-921 + addLayer :806-894)
    def add_layer(self, weight, bias, activation="None", activation_param=1.0):
        w = np.asarray(weight, dtype=np.float32).astype(np.float16)
        b = np.asarray(bias, dtype=np.float32).astype(np.float16)
        co, ci = w.shape
        if not self.layers and self.num_fourier > 0:
            if not self.has_time:
                if self.has_direction:
                    wn = np.zeros((co, ci + 2), np.float16)
                    wn[:, 0:3] = w[:, 0:3]; wn[:, 4:7] = w[:, 3:6]; wn[:, 8:] = w[:, 6:]
                else:
                    wn = np.zeros((co, ci + 1), np.float16)
                    wn[:, 0:3] = w[:, 0:3]; wn[:, 4:] = w[:, 3:]
                w = wn
            elif self.has_direction:
                wn = np.zeros((co, ci + 1), np.float16)
                wn[:, 0:4] = w[:, 0:4]; wn[:, 4:7] = w[:, 4:7]; wn[:, 8:] = w[:, 7:]
                w = wn
        self.layers.append(Layer(w, b, activation, float(activation_param)))

    def set_fourier_matrix(self, B, premultiplied: bool):
        B = np.asarray(B, dtype=np.float32)
        assert B.ndim == 2 and B.shape[1] in (3, 6)
        self.direction_in_fourier = B.shape[1] == 6
        f = (1.0 if premultiplied else 2.0 * np.pi) * B.astype(np.float64)
        self.fourier = f.astype(np.float32).astype(np.float16)

    def valid(self):
        cur = self.input_channels_out() + self.latent_channels()
        for i, l in enumerate(self.layers):
            if l.cin != cur:
                return False
            cur = l.cout
        return cur == OUTPUT_CHANNELS_IN[OUTPUT_MODES.index(self.output_mode)] and self.num_fourier % 2 == 0


# ------------------------------------------------------------------------------------------------ reading
class _Reader:
    def __init__(self, data: bytes):
        self.d = memoryview(data); self.p = 0

    def take(self, n):
        if self.p + n > len(self.d):
            raise ValueError("volnet: unexpected end of data")
        v = self.d[self.p:self.p + n]; self.p += n
        return v

    def i32(self): return struct.unpack("<i", self.take(4))[0]
    def u8(self): return struct.unpack("<B", self.take(1))[0]
    def f32(self): return struct.unpack("<f", self.take(4))[0]
    def string(self): return bytes(self.take(self.i32())).decode("ascii")
    def array(self, dtype, count): return np.frombuffer(self.take(np.dtype(dtype).itemsize * count), dtype=dtype).copy()


def _erfinv(x):
    # Winitzki initial guess + Newton iterations on erf (float64)
    from math import erf, pi
    x = np.asarray(x, dtype=np.float64)
    a = 0.147
    ln = np.log(np.maximum(1 - x * x, 1e-300))
    t = 2 / (np.pi * a) + ln / 2
    y = np.sign(x) * np.sqrt(np.sqrt(t * t - ln / a) - t)
    verf = np.vectorize(erf)
    for _ in range(3):
        y = y - (verf(y) - x) / (2 / np.sqrt(pi) * np.exp(-y * y))
    return y


def _grid_from_storage(raw_flat: np.ndarray, C, Z, Y, X):
    """storage order (C/4, Z, Y, X, 4) -> dense (C,Z,Y,X)  (volume_interpolation_network.h:243-246)"""
    return raw_flat.reshape(C // 4, Z, Y, X, 4).transpose(0, 4, 1, 2, 3).reshape(C, Z, Y, X)


def _grid_to_storage(dense: np.ndarray):
    C, Z, Y, X = dense.shape
    return np.ascontiguousarray(dense.reshape(C // 4, 4, Z, Y, X).transpose(0, 2, 3, 4, 1)).reshape(-1)


def _load_grid(r: _Reader) -> LatentGrid:
    version = r.i32()
    if version != 1:
        raise ValueError(f"Unknown version for LatentGrid {version}")
    enc, C, Z, Y, X = r.i32(), r.i32(), r.i32(), r.i32(), r.i32()
    n = C * Z * Y * X
    if enc == GRID_FLOAT:
        vals = _grid_from_storage(r.array("<f4", n), C, Z, Y, X).astype(np.float32)
        return LatentGrid(enc, np.ascontiguousarray(vals))
    raw = _grid_from_storage(r.array("u1", n), C, Z, Y, X)
    off = r.array("<f4", C); sc = r.array("<f4", C)
    u = raw.astype(np.float32) / np.float32(255.0)
    if enc == GRID_BYTE_LINEAR:
        vals = off[:, None, None, None] + u * sc[:, None, None, None]
    else:
        t = (np.sqrt(2.0) * _erfinv((np.float32(2.0) - np.float32(1e-4)) * (u - np.float32(0.5)))).astype(np.float32)
        vals = off[:, None, None, None] + t * sc[:, None, None, None]
    return LatentGrid(enc, np.ascontiguousarray(vals.astype(np.float32)), np.ascontiguousarray(raw), off, sc)


def loads(data: bytes) -> SceneNetwork:
    r = _Reader(data)
    net = SceneNetwork()
    version = r.i32()
    if version not in (1, 2):
        raise ValueError(f"Unknown version for SceneNetwork {version}")
    v = r.i32()
    if v not in (1, 2, 3):
        raise ValueError(f"Unknown version for InputParametrization {v}")
    if v == 3:
        net.has_time = r.u8() != 0
    net.has_direction = r.u8() != 0
    nf = r.i32()
    net.direction_in_fourier = (r.u8() != 0) if v >= 2 else False
    fc = 6 if net.direction_in_fourier else 3
    if nf > 0:
        net.fourier = np.ascontiguousarray(r.array("<f2", nf * fc).reshape(fc, nf).T)  # stored index = cout + nf*cin
    v = r.i32()
    if v != 1:
        raise ValueError(f"Unknown version for OutputParametrization {v}")
    net.output_mode = r.string()
    if net.output_mode not in OUTPUT_MODES:
        raise ValueError("No output mode found matching string " + net.output_mode)
    num_layers = r.i32()
    for i in range(num_layers):
        v = r.i32()
        if v not in (1, 2):
            raise ValueError(f"Unknown version for Layer {v}")
        rows, cols = r.i32(), r.i32()
        w = r.array("<f2", rows * cols); b = r.array("<f2", rows)
        act = r.string()
        if act not in ACTIVATIONS:
            raise ValueError("No activation found matching string " + act)
        param = r.f32() if v == 2 else 1.0
        transposed = not (i == 0 and nf > 0) and (cols < 16 or rows < 16)
        w = w.reshape(cols, rows).T if transposed else w.reshape(rows, cols)
        net.layers.append(Layer(np.ascontiguousarray(w), b, act, param))
    net.box_min = tuple(float(x) for x in r.array("<f4", 3))
    net.box_size = tuple(float(x) for x in r.array("<f4", 3))
    if version == 2 and r.u8() > 0:
        v = r.i32()
        if v > 1:
            raise ValueError(f"Unknown version for LatentGridTimeAndEnsemble {v}")
        net.has_latent_grid = True
        net.time_min, net.time_num, net.time_step = r.i32(), r.i32(), r.i32()
        net.ensemble_min, net.ensemble_num = r.i32(), r.i32()
        net.time_grids = [_load_grid(r) for _ in range(net.time_num)]
        net.ensemble_grids = [_load_grid(r) for _ in range(net.ensemble_num)]
    return net


# ------------------------------------------------------------------------------------------------ writing
def _save_grid(out: list, g: LatentGrid):
    C, Z, Y, X = g.shape
    out.append(struct.pack("<6i", 1, g.encoding, C, Z, Y, X))
    if g.encoding == GRID_FLOAT:
        out.append(_grid_to_storage(g.values.astype("<f4")).tobytes())
    else:
        out.append(_grid_to_storage(g.raw.astype("u1")).tobytes())
        out.append(np.asarray(g.offset_or_mean, "<f4").tobytes()); out.append(np.asarray(g.scale_or_std, "<f4").tobytes())


def dumps(net: SceneNetwork) -> bytes:
    out = [struct.pack("<i", 2)]
    nf = net.num_fourier
    out.append(struct.pack("<iBBiB", 3, int(net.has_time), int(net.has_direction), nf, int(net.direction_in_fourier)))
    if nf > 0:
        out.append(np.ascontiguousarray(net.fourier.astype("<f2").T).tobytes())
    mode = net.output_mode.encode("ascii")
    out.append(struct.pack("<ii", 1, len(mode)) + mode)
    out.append(struct.pack("<i", len(net.layers)))
    for i, l in enumerate(net.layers):
        rows, cols = l.cout, l.cin
        transposed = not (i == 0 and nf > 0) and (cols < 16 or rows < 16)
        w = l.weights.astype("<f2")
        wb = np.ascontiguousarray(w.T).tobytes() if transposed else np.ascontiguousarray(w).tobytes()
        act = l.activation.encode("ascii")
        out.append(struct.pack("<iii", 2, rows, cols) + wb + l.bias.astype("<f2").tobytes()
                   + struct.pack("<i", len(act)) + act + struct.pack("<f", l.activation_param))
    out.append(struct.pack("<3f", *net.box_min) + struct.pack("<3f", *net.box_size))
    out.append(struct.pack("<B", 1 if net.has_latent_grid else 0))
    if net.has_latent_grid:
        out.append(struct.pack("<6i", 1, net.time_min, net.time_num, net.time_step, net.ensemble_min, net.ensemble_num))
        for g in net.time_grids: _save_grid(out, g)
        for g in net.ensemble_grids: _save_grid(out, g)
    return b"".join(out)


def load(path) -> SceneNetwork:
    with open(path, "rb") as f:
        return loads(f.read())


def save(net: SceneNetwork, path):
    with open(path, "wb") as f:
        f.write(dumps(net))
