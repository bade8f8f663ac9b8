"""
Host-side mirror of the reference's ``pyrenderer`` module for the SRN/DVR hot path.

Same class names, attribute names, argument meaning and error behaviour (RuntimeError) as the pybind11 module
built from bindings/bindings.cpp + the ``registerPybindModule`` functions of
renderer/volume_interpolation_network.cpp:1817-2007, renderer/ray_evaluation_stepping.cpp:781-801,
renderer/image_evaluator_simple.cpp:427-475, renderer/iimage_evaluator.cpp:325-358,
renderer/transfer_function_piecewise.cpp:110-111, renderer/transfer_function_texture.cpp:246-264 and
renderer/camera.cpp:200-222,385-396 — so that callers written against the reference
(applications/volnet/inference.py:461-625) run unchanged. All compute goes through the C ABI of
libfvsrn_b200.so; torch tensors exist only at this boundary (device memory + current stream).
"""
from __future__ import annotations

import ctypes as C
import math
import os
from typing import Optional

import numpy as np
import torch

from . import _lib
from ._lib import check


# Kernel family used by networks created from now on (B200 addition, see fvsrn_network_set_kernel_mode):
# "tcgen05" (default; tcgen05.mma + TMEM, fp32 accumulate) or "mma_sync" (reference-precision fp16-accumulate mode).
KERNEL_MODES = {"mma_sync": 0, "tcgen05": 1, "reference_exact": 2}
DEFAULT_KERNEL_MODE = __import__("os").environ.get("FVSRN_KERNEL_MODE", "tcgen05")


# ------------------------------------------------------------------------------------------------ small value types
class _Vec3:
    __slots__ = ("x", "y", "z")

    def __init__(self, x=0.0, y=0.0, z=0.0):
        self.x, self.y, self.z = float(x), float(y), float(z)

    def __iter__(self): return iter((self.x, self.y, self.z))
    def __repr__(self): return f"{type(self).__name__}({self.x}, {self.y}, {self.z})"


class float3(_Vec3): pass
class double3(_Vec3): pass


class _Vec4:
    __slots__ = ("x", "y", "z", "w")

    def __init__(self, x=0.0, y=0.0, z=0.0, w=0.0):
        self.x, self.y, self.z, self.w = float(x), float(y), float(z), float(w)

    def __iter__(self): return iter((self.x, self.y, self.z, self.w))


class float4(_Vec4): pass
class double4(_Vec4): pass


class Parameter:
    """renderer/parameter.h:19-107: ``.value`` is a scalar/vector or a batched torch tensor."""

    def __init__(self, value=None):
        self.value = value
        self.grad = None
        self.forwardIndex = None


def _stream_ptr():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def sync():
    torch.cuda.synchronize()


def cleanup():
    pass


class GPUTimer:
    """cudaEvent pair, bindings/bindings.cpp:101-131."""

    def __init__(self):
        self._a = torch.cuda.Event(enable_timing=True); self._b = torch.cuda.Event(enable_timing=True)

    def start(self): self._a.record()
    def stop(self): self._b.record()

    def elapsed_milliseconds(self):
        self._b.synchronize()
        return self._a.elapsed_time(self._b)


# ------------------------------------------------------------------------------------------------ SceneNetwork
class _Enum:
    def __init__(self, name, index): self.name, self.index = name, index
    def __repr__(self): return self.name
    def __eq__(self, o): return isinstance(o, _Enum) and o.index == self.index and type(o) is type(self)
    def __hash__(self): return hash((type(self).__name__, self.index))


class SceneNetwork:
    """renderer::SceneNetwork (volume_interpolation_network.cpp:797-1113, pybind :1828-1972)."""

    class Layer:
        class Activation(_Enum): pass
        _names = ["ReLU", "Sine", "Snake", "SnakeAlt", "Sigmoid", "None"]

        @staticmethod
        def ActivationFromString(s):
            if s not in SceneNetwork.Layer._names:
                raise RuntimeError("No output mode found matching string " + s)
            return SceneNetwork.Layer.Activation(s, SceneNetwork.Layer._names.index(s))

    class OutputParametrization:
        class OutputMode(_Enum): pass
        _names = ["density", "density:direct", "rgbo", "rgbo:direct", "densitygrad", "densitygrad:direct",
                  "densitygrad:cubic", "densitycurvature", "densitycurvature:direct"]

        @staticmethod
        def OutputModeFromString(s):
            names = SceneNetwork.OutputParametrization._names
            if s not in names:
                raise RuntimeError("No output mode found matching string " + s)
            return SceneNetwork.OutputParametrization.OutputMode(s, names.index(s))

        def __init__(self, owner): self._o = owner

        @property
        def output_mode(self):
            return SceneNetwork.OutputParametrization.OutputMode(
                SceneNetwork.OutputParametrization._names[self._o._info().output_mode], self._o._info().output_mode)

        @output_mode.setter
        def output_mode(self, m):
            name = m.name if isinstance(m, _Enum) else str(m)
            check(_lib.lib().fvsrn_network_set_output_mode(self._o._h, name.encode()))

    class LatentGrid:
        class Encoding(_Enum): pass
    LatentGrid.Float = LatentGrid.Encoding("Float", 0)
    LatentGrid.ByteLinear = LatentGrid.Encoding("ByteLinear", 1)
    LatentGrid.ByteGaussian = LatentGrid.Encoding("ByteGaussian", 2)

    class LatentGridTimeAndEnsemble:
        """(time_min, time_num, time_step, ensemble_min, ensemble_num), :756-796."""

        def __init__(self, time_min=0, time_num=0, time_step=1, ensemble_min=0, ensemble_num=0):
            self.time_min, self.time_num, self.time_step = int(time_min), int(time_num), int(time_step)
            self.ensemble_min, self.ensemble_num = int(ensemble_min), int(ensemble_num)
            self._time = [None] * self.time_num
            self._ens = [None] * self.ensemble_num

        @staticmethod
        def _dense(t):
            if t.dim() != 5 or t.shape[0] != 1:
                raise RuntimeError("Expected a tensor of shape (1,C,Z,Y,X)")
            return np.ascontiguousarray(t.detach().to("cpu", torch.float32).numpy()[0])

        def set_time_grid_from_torch(self, index, t, encoding):
            if not (0 <= index < self.time_num):
                raise RuntimeError("index out of bounds!")
            self._time[index] = (self._dense(t), encoding.index)
            return 0.0

        def set_ensemble_grid_from_torch(self, index, t, encoding):
            if not (0 <= index < self.ensemble_num):
                raise RuntimeError("index out of bounds!")
            self._ens[index] = (self._dense(t), encoding.index)
            return 0.0

        def is_valid(self):
            gs = self._time + self._ens
            return bool(gs) and all(g is not None and g[0].shape[0] % 16 == 0 for g in gs)

    class _Input:
        """InputParametrization proxy (:68-156)."""

        def __init__(self, owner):
            self._o = owner; self._has_time = False; self._has_direction = False

        def _push(self): check(_lib.lib().fvsrn_network_set_input(self._o._h, int(self._has_time), int(self._has_direction)))

        @property
        def has_time(self): return bool(self._o._info().has_time)

        @has_time.setter
        def has_time(self, v): self._has_time = bool(v); self._push()

        @property
        def has_direction(self): return bool(self._o._info().has_direction)

        @has_direction.setter
        def has_direction(self, v): self._has_direction = bool(v); self._push()

        @property
        def num_fourier_features(self): return self._o._info().num_fourier

        def set_fourier_matrix_from_tensor(self, t, premultiplied):
            if t.dim() != 2:
                raise RuntimeError("Expected a 2D tensor")
            a = np.ascontiguousarray(t.detach().to("cpu", torch.float32).numpy())
            check(_lib.lib().fvsrn_network_set_fourier(self._o._h, a.ctypes.data, a.shape[0], a.shape[1], int(bool(premultiplied))))

        def disable_fourier_features(self):
            check(_lib.lib().fvsrn_network_disable_fourier(self._o._h))

    def __init__(self, _handle=None):
        L = _lib.lib()
        if _handle is None:
            h = C.c_void_p()
            check(L.fvsrn_network_create(C.byref(h)))
            _handle = h
        self._h = _handle
        self.input = SceneNetwork._Input(self)
        if _handle is not None:
            i = self._info(); self.input._has_time, self.input._has_direction = bool(i.has_time), bool(i.has_direction)
        self.output = SceneNetwork.OutputParametrization(self)
        self._latent_grid = None
        self._prepared_device = None
        self.kernel_mode = DEFAULT_KERNEL_MODE

    def __del__(self):
        try:
            if getattr(self, "_h", None):
                _lib.lib().fvsrn_network_free(self._h); self._h = None
        except Exception:
            pass

    def _info(self):
        info = _lib.NetworkInfo()
        check(_lib.lib().fvsrn_network_get_info(self._h, C.byref(info)))
        return info

    # -- latent grid
    @property
    def latent_grid(self): return self._latent_grid

    @latent_grid.setter
    def latent_grid(self, g):
        L = _lib.lib()
        check(L.fvsrn_network_set_latent_grid_layout(self._h, g.time_min, g.time_num, g.time_step, g.ensemble_min, g.ensemble_num))
        for idx, item in enumerate(g._time):
            if item is None:
                raise RuntimeError("One latent grid was null")
            a, enc = item
            check(L.fvsrn_network_set_time_grid(self._h, idx, a.ctypes.data, *a.shape, enc, None))
        for idx, item in enumerate(g._ens):
            if item is None:
                raise RuntimeError("One latent grid was null")
            a, enc = item
            check(L.fvsrn_network_set_ensemble_grid(self._h, idx, a.ctypes.data, *a.shape, enc, None))
        self._latent_grid = g
        self._prepared_device = None

    # -- layers
    def add_layer(self, weights, bias, activation, activation_parameter=1.0):
        if weights.dim() != 2 or bias.dim() != 1 or weights.shape[0] != bias.shape[0]:
            raise RuntimeError("Expected weights (cout,cin) and bias (cout)")
        w = np.ascontiguousarray(weights.detach().to("cpu", torch.float32).numpy())
        b = np.ascontiguousarray(bias.detach().to("cpu", torch.float32).numpy())
        name = activation.name if isinstance(activation, _Enum) else str(activation)
        check(_lib.lib().fvsrn_network_add_layer(self._h, w.ctypes.data, b.ctypes.data, w.shape[0], w.shape[1],
                                                 name.encode(), float(activation_parameter)))
        self._prepared_device = None

    # -- box
    def _get_box(self):
        mn, sz = (C.c_float * 3)(), (C.c_float * 3)()
        check(_lib.lib().fvsrn_network_get_box(self._h, C.byref(mn), C.byref(sz)))
        return float3(*mn), float3(*sz)

    def _set_box(self, mn, sz):
        check(_lib.lib().fvsrn_network_set_box(self._h, C.byref((C.c_float * 3)(*mn)), C.byref((C.c_float * 3)(*sz))))

    @property
    def box_min(self): return self._get_box()[0]

    @box_min.setter
    def box_min(self, v): self._set_box(tuple(v), tuple(self._get_box()[1]))

    @property
    def box_size(self): return self._get_box()[1]

    @box_size.setter
    def box_size(self, v): self._set_box(tuple(self._get_box()[0]), tuple(v))

    def valid(self): return _lib.lib().fvsrn_network_valid(self._h) == 1
    def num_parameters(self): return _lib.lib().fvsrn_network_num_parameters(self._h)
    def compute_max_warps(self, only_shared, adjoint=False):
        return _lib.lib().fvsrn_network_compute_max_warps(self._h, int(bool(only_shared)), int(bool(adjoint)))

    def clear_gpu_resources(self):
        check(_lib.lib().fvsrn_network_clear_gpu_resources(self._h)); self._prepared_device = None

    def set_time_and_ensemble(self, time, ensemble):
        check(_lib.lib().fvsrn_network_set_time_and_ensemble(self._h, float(time), int(ensemble)))

    def to_bytes(self) -> bytes:
        n = C.c_size_t()
        check(_lib.lib().fvsrn_network_save_volnet(self._h, None, 0, C.byref(n)))
        buf = C.create_string_buffer(n.value)
        check(_lib.lib().fvsrn_network_save_volnet(self._h, buf, n.value, C.byref(n)))
        return buf.raw[:n.value]

    def save(self, path):
        data = self.to_bytes()
        with open(path, "wb") as f:
            f.write(data)

    @staticmethod
    def from_bytes(data: bytes):
        h = C.c_void_p()
        check(_lib.lib().fvsrn_network_load_volnet(data, len(data), C.byref(h)))
        return SceneNetwork(h)

    @staticmethod
    def load(path):
        with open(path, "rb") as f:
            return SceneNetwork.from_bytes(f.read())

    # -- GPU residency (replaces fillConstantMemory / copyGridToGPU)
    def _ensure_prepared(self):
        dev = torch.cuda.current_device() if torch.cuda.is_available() else 0
        if self.kernel_mode not in KERNEL_MODES:
            raise RuntimeError(f"unknown kernel mode {self.kernel_mode!r} (expected one of {sorted(KERNEL_MODES)})")
        check(_lib.lib().fvsrn_network_set_kernel_mode(self._h, KERNEL_MODES[self.kernel_mode]))
        if self._prepared_device != dev:
            check(_lib.lib().fvsrn_network_prepare(self._h, dev, _stream_ptr() if torch.cuda.is_available() else None))
            self._prepared_device = dev


# ------------------------------------------------------------------------------------------------ volume module
class VolumeInterpolationNetwork:
    """renderer::VolumeInterpolationNetwork (:1419-1726, pybind :1974-2005) + IVolumeInterpolation::evaluate."""

    class GradientMode(_Enum): pass
    OFF_OR_DIRECT = GradientMode("OFF_OR_DIRECT", 0)             # no normals, or the ones a densitygrad* network predicts
    FINITE_DIFFERENCES = GradientMode("FINITE_DIFFERENCES", 1)   # 6 more SRN evaluations per shaded step
    ADJOINT_METHOD = GradientMode("ADJOINT_METHOD", 2)           # one backward pass through the network per shaded step

    def __init__(self):
        self._networks = []
        self._current = None
        self.only_shared_memory = False   # kept for API compatibility; the B200 kernel always pins weights in smem
        self.gradient_mode = VolumeInterpolationNetwork.OFF_OR_DIRECT
        self.finite_differences_stepsize = 1.0 / 256.0     # volume_interpolation_network.cpp:1423
        self.adjoint_latent_grid_central_differences_stepsize_scale = 4.0     # :1424, pybind :1997-2004

    def _apply_gradient_mode(self, net: "SceneNetwork"):
        """Pushes the module's gradient settings to the network handle (the reference passes them as defines / constants per render)."""
        gm = self.gradient_mode
        index = gm.index if isinstance(gm, _Enum) else {"OFF_OR_DIRECT": 0, "FINITE_DIFFERENCES": 1, "ADJOINT_METHOD": 2}[str(gm)]
        step = float(self.finite_differences_stepsize)
        if index == 2:      # fillConstantMemory, :1809-1812: 1 / (float(gridResolution) * scale), computed in float
            res = int(net._info().grid_x)
            step = float(np.float32(1.0) / (np.float32(res) * np.float32(self.adjoint_latent_grid_central_differences_stepsize_scale))) if res > 0 else 0.0
        check(_lib.lib().fvsrn_network_set_gradient_mode(net._h, index, step))

    def set_network(self, net: SceneNetwork):
        if net is None or not net.valid():
            raise RuntimeError("scene network is not valid")
        self._networks = [net]; self._current = net

    def add_network(self, net: SceneNetwork):
        if net is None or not net.valid():
            raise RuntimeError("scene network is not valid")
        self._networks.append(net)
        if self._current is None:
            self._current = net

    def select_network(self, index): self._current = self._networks[index]
    def load_network(self, path): self.set_network(SceneNetwork.load(path))
    def current_network(self): return self._current
    def box_min(self): return self._require().box_min
    def box_size(self): return self._require().box_size

    def _require(self) -> SceneNetwork:
        if self._current is None:
            raise RuntimeError("No network specified")
        return self._current

    def evaluate(self, positions: torch.Tensor, directions: Optional[torch.Tensor] = None) -> torch.Tensor:
        """positions (N,3) CUDA float32 in [0,1]^3 -> (N,C) float32 (volume_interpolation.cpp:26-127)."""
        net = self._require()
        if not positions.is_cuda:
            raise RuntimeError("positions must reside on the GPU")
        if positions.dim() != 2 or positions.shape[1] != 3:
            raise RuntimeError("positions must be of shape (N,3)")
        if positions.dtype != torch.float32:
            raise RuntimeError("positions must be of dtype float32")
        positions = positions.contiguous()
        if directions is not None:
            directions = directions.to(torch.float32).contiguous()
        net._ensure_prepared()
        cout = net._info().channels_out
        out = torch.empty((positions.shape[0], cout), dtype=torch.float32, device=positions.device)
        check(_lib.lib().fvsrn_evaluate(net._h, positions.data_ptr(), directions.data_ptr() if directions is not None else None,
                                        out.data_ptr(), positions.shape[0], _stream_ptr()))
        return out


# ------------------------------------------------------------------------------------------------ ground-truth volume
class Volume:
    """renderer::Volume (volume.h:32-77): the .cvol container. `Volume(path)` loads, `Volume.from_numpy` wraps a density grid."""

    def __init__(self, path: Optional[str] = None, _handle=None):
        self._h = _handle
        if path is not None:
            with open(path, "rb") as f:
                data = f.read()
            h = C.c_void_p()
            check(_lib.lib().fvsrn_volume_load_cvol(data, len(data), C.byref(h)))
            self._h = h
        self._prepared = None

    @staticmethod
    def from_bytes(data: bytes) -> "Volume":
        h = C.c_void_p()
        check(_lib.lib().fvsrn_volume_load_cvol(data, len(data), C.byref(h)))
        return Volume(_handle=h)

    @staticmethod
    def from_numpy(density: np.ndarray, world_size=(1.0, 1.0, 1.0)) -> "Volume":
        """density: (Z,Y,X) float32, x fastest (Volume::MipmapLevel::idx, volume.h:126-132)."""
        a = np.ascontiguousarray(density, np.float32)
        if a.ndim != 3:
            raise RuntimeError("expected a (Z,Y,X) density grid")
        h = C.c_void_p()
        check(_lib.lib().fvsrn_volume_from_dense(a.ctypes.data, a.shape[2], a.shape[1], a.shape[0], C.byref((C.c_float * 3)(*world_size)), C.byref(h)))
        return Volume(_handle=h)

    def __del__(self):
        try:
            if getattr(self, "_h", None):
                _lib.lib().fvsrn_volume_free(self._h); self._h = None
        except Exception:
            pass

    def info(self, feature=0):
        i = _lib.VolumeInfo()
        check(_lib.lib().fvsrn_volume_get_info(self._h, feature, C.byref(i)))
        return i

    def num_features(self): return self.info().num_features
    def world_size(self): return float3(*self.info().world_size)

    def to_bytes(self) -> bytes:
        n = C.c_size_t()
        check(_lib.lib().fvsrn_volume_save_cvol(self._h, None, 0, C.byref(n)))
        buf = C.create_string_buffer(n.value)
        check(_lib.lib().fvsrn_volume_save_cvol(self._h, buf, n.value, C.byref(n)))
        return buf.raw[:n.value]

    def save(self, path, compression=0):
        if compression:
            raise RuntimeError("the B200 path writes uncompressed .cvol files")
        with open(path, "wb") as f:
            f.write(self.to_bytes())

    def _ensure_prepared(self, feature=0):
        dev = torch.cuda.current_device() if torch.cuda.is_available() else 0
        if self._prepared != (dev, feature):
            check(_lib.lib().fvsrn_volume_prepare(self._h, dev, feature, _stream_ptr() if torch.cuda.is_available() else None))
            self._prepared = (dev, feature)


class VolumeInterpolationGrid:
    """renderer::VolumeInterpolationGrid with a Volume source, trilinear interpolation (volume_interpolation_grid.cpp:153-198,
    renderer_volume_grid.cuh): the ground truth the SRN renders are compared with. Assign it to ImageEvaluatorSimple.volume."""

    def __init__(self):
        self._volume: Optional[Volume] = None
        self.feature = 0

    def setSource(self, volume: Volume, mipmap: int = 0):
        if mipmap != 0:
            raise RuntimeError("mipmap levels are not part of the B200 path")
        self._volume = volume

    set_source = setSource

    def volume(self): return self._volume
    def mipmap_level(self): return 0

    def _require(self) -> Volume:
        if self._volume is None:
            raise RuntimeError("No volume specified, can't render!")
        return self._volume

    def box_min(self): return float3(*self._require().info(self.feature).box_min)
    def box_size(self): return float3(*self._require().info(self.feature).box_size)

    def evaluate(self, positions: torch.Tensor, directions=None) -> torch.Tensor:
        v = self._require()
        if not positions.is_cuda or positions.dim() != 2 or positions.shape[1] != 3 or positions.dtype != torch.float32:
            raise RuntimeError("positions must be a CUDA float32 tensor of shape (N,3)")
        positions = positions.contiguous()
        v._ensure_prepared(self.feature)
        out = torch.empty((positions.shape[0], 1), dtype=torch.float32, device=positions.device)
        check(_lib.lib().fvsrn_evaluate_grid(v._h, positions.data_ptr(), out.data_ptr(), positions.shape[0], _stream_ptr()))
        return out


# ------------------------------------------------------------------------------------------------ TF modules
class ITransferFunction:
    kind = -1

    def _tensor(self) -> torch.Tensor: raise NotImplementedError

    def _evaluate(self, density, min_density, max_density, stepsize):
        if not density.is_cuda:
            raise RuntimeError("density must reside on the GPU")
        if density.dim() != 2 or density.shape[1] != 1:
            raise RuntimeError("density must be of shape (N,1)")
        d = density.to(torch.float32).contiguous()
        tf = self._tensor()
        out = torch.empty((d.shape[0], 4), dtype=torch.float32, device=d.device)
        check(_lib.lib().fvsrn_evaluate_tf(self.kind, tf.data_ptr(), tf.shape[1] if self.kind != 2 else 0, d.data_ptr(), out.data_ptr(),
                                           d.shape[0], float(min_density), float(max_density), float(stepsize), _stream_ptr()))
        return out

    def evaluate(self, density, min_density, max_density):
        """ITransferFunction::evaluate (transfer_function.cpp:123-136 -> EvaluateTF, renderer_tf_kernels.cuh:11-40):
        density (N,1) -> rgb + absorption (N,4), stepsize 1."""
        return self._evaluate(density, min_density, max_density, 1.0)

    def evaluate_with_previous(self, density, min_density, max_density, previous_density, stepsize):
        """ITransferFunction::evaluate_with_previous (EvaluateTFWithPrevious, :42-70); the previous density only enters
        pre-integrated texture TFs, which are not on the path."""
        return self._evaluate(density, min_density, max_density, stepsize)


class TransferFunctionPiecewiseLinear(ITransferFunction):
    """tensor (B,R,5) = r,g,b,absorption,position (transfer_function_piecewise.cpp:145-164, 264-278)."""
    kind = 0

    def __init__(self):
        self.tensor = Parameter(None)
        self._absorption_scaling = 1.0
        self._cpu: Optional[np.ndarray] = None

    def load_points(self, color_points, opacity_points, absorption_scaling: float):
        """TransferFunctionPiecewiseLinear::load + computeTensor (transfer_function_piecewise.cpp:80-91, 166-279): the colorPoints /
        opacityPoints / absorptionScaling of a settings file -> `tensor`."""
        from . import settings
        self._cpu = settings.piecewise_tf_tensor(color_points, opacity_points, absorption_scaling)
        self._absorption_scaling = float(absorption_scaling)
        self.tensor = Parameter(torch.from_numpy(self._cpu).to("cuda"))

    def evaluate_scalar(self, density: float):
        """TransferFunctionPiecewiseLinear::evaluate(double) (:281-285): rgba at one density from the CPU copy of the tensor."""
        from . import settings
        t = self._cpu if self._cpu is not None else self._tensor().cpu().numpy()
        return double4(*(float(v) for v in settings.sample_piecewise(t, density)))

    def get_max_absorption(self): return self._absorption_scaling

    def _tensor(self):
        t = self.tensor.value
        if t is None:
            raise RuntimeError("texture tensor is not defined")
        if not t.is_cuda:
            raise RuntimeError("texture tensor must reside on the GPU")
        if t.dim() != 3 or t.shape[2] != 5:
            raise RuntimeError("texture tensor must be of shape (B,R,5)")
        return t.to(torch.float32).contiguous()


class TransferFunctionTexture(ITransferFunction):
    """tensor (B,R,4) rgba, sampled like a clamped, linearly filtered 1-D texture (renderer_tf_texture.cuh:46-56)."""
    kind = 1

    def __init__(self):
        self.tensor = Parameter(None)
        self.preintegration_mode = 0
        self._absorption_scaling = 10.0          # transfer_function_texture.cpp:17

    def switch_to_tensor(self): pass

    def load_points(self, color_points, plot, absorption_scaling: float):
        """TransferFunctionTexture::load + computeTexture (transfer_function_texture.cpp:215-226, 347-362): colorPoints, the 256
        opacityPoints and absorptionScaling of a settings file -> the (1,256,4) texture."""
        from . import settings
        self._absorption_scaling = float(absorption_scaling)
        self.tensor = Parameter(torch.from_numpy(settings.texture_tf_tensor(color_points, plot, absorption_scaling)).to("cuda"))

    def paste(self, other: "ITransferFunction"):
        """TransferFunctionTexture::doPaste (:412-437): resample another TF at the 256 texel centres."""
        from . import settings
        if not hasattr(other, "evaluate_scalar"):
            raise RuntimeError("Copying to texture TF not supported from source TF")
        cps, plot, scale = settings.paste_into_texture(lambda d: tuple(other.evaluate_scalar(d)), self._absorption_scaling)
        self.load_points(cps, plot, scale)

    def get_max_absorption(self): return self._absorption_scaling

    def _tensor(self):
        if self.preintegration_mode != 0:
            raise RuntimeError("pre-integrated texture TFs are not part of the B200 hot path")
        t = self.tensor.value
        if t is None:
            raise RuntimeError("texture tensor is not defined")
        if not t.is_cuda or t.dim() != 3 or t.shape[2] != 4:
            raise RuntimeError("texture tensor must be a CUDA tensor of shape (B,R,4)")
        return t.to(torch.float32).contiguous()


class TransferFunctionIdentity(ITransferFunction):
    """rgb = density*scale_emission, absorption = density*scale_absorption (renderer_tf_identity.cuh)."""
    kind = 2

    def __init__(self, scale_absorption=10.0, scale_emission=1.0):
        self.absorption_scaling = float(scale_absorption)
        self.emission_scaling = float(scale_emission)

    def _tensor(self):
        return torch.tensor([[[self.absorption_scaling, self.emission_scaling]]], dtype=torch.float32, device="cuda")

    def evaluate_scalar(self, density: float):
        return double4(density * self.emission_scaling, density * self.emission_scaling, density * self.emission_scaling, density * self.absorption_scaling)

    def get_max_absorption(self): return self.absorption_scaling


class Blending:
    class BlendMode(_Enum): pass
    Alpha = BlendMode("Alpha", 0)
    BeerLambert = BlendMode("BeerLambert", 1)

    def __init__(self):
        self.blendMode = Blending.BeerLambert   # reference default, blending.h:54


class BRDFLambert:
    """renderer::BRDFLambert with scalar parameters (brdf.cpp:124-135, 204-225, pybind :256-274; kernel renderer_brdf_lambert.cuh:57-103).
    Both options need normals: a densitygrad* network or VolumeInterpolationNetwork.gradient_mode = FINITE_DIFFERENCES."""

    class LightType(_Enum): pass
    Point = LightType("Point", 0)
    Directional = LightType("Directional", 1)

    def __init__(self):
        self.enable_magnitude_scaling = False
        self.enable_phong = False
        self.magnitude_scaling = 1.0
        self.ambient = 1.0
        self.specular = 1.0
        self.magnitude_center = 1.0
        self.magnitude_radius = 1.0
        self.specular_exponent = 1
        self.light_follows_camera = True
        self.light_type = BRDFLambert.Directional
        self.light_position = double3(0, 0, 0)
        self.light_direction = double3(0, 0, 0)

    def _fill(self, p, cam: torch.Tensor):
        flags = (1 if self.enable_magnitude_scaling else 0) | (2 if self.enable_phong else 0)
        if not flags:
            return
        pos, direction = tuple(self.light_position), tuple(self.light_direction)
        if self.light_follows_camera:      # updateLightFromCamera, brdf.cpp:478-497: origin and front of camera 0
            c = cam[0].detach().cpu().numpy().astype(np.float64)
            pos = tuple(c[0])
            direction = tuple(np.cross(c[2], c[1]))     # front = cross(up, right), renderer_camera.cuh:38
        p.brdf_flags, p.light_type = flags, self.light_type.index
        p.light = (C.c_float * 3)(*(pos if self.light_type.index == 0 else direction))
        p.brdf_magnitude_scaling, p.brdf_ambient, p.brdf_specular = float(self.magnitude_scaling), float(self.ambient), float(self.specular)
        p.brdf_magnitude_center, p.brdf_magnitude_radius = float(self.magnitude_center), float(self.magnitude_radius)
        p.brdf_specular_exponent = int(self.specular_exponent)


class RayEvaluationSteppingDvr:
    """renderer::RayEvaluationSteppingDvr (ray_evaluation_stepping.cpp:535-623, pybind :781-801)."""

    def __init__(self):
        self.stepsize = 0.005            # ray_evaluation_stepping.h:14
        self.min_density = 0.0
        self.max_density = 1.0
        self.early_out = True
        self.alpha_early_out = 1.0 - 1e-5
        self.blending = Blending()
        self.tf: Optional[ITransferFunction] = TransferFunctionPiecewiseLinear()
        self.brdf = BRDFLambert()
        self.volume = None

    def convert_to_texture_tf(self):
        """RayEvaluationSteppingDvr::convertToTextureTF (ray_evaluation_stepping.cpp:757-770)."""
        if self.tf is None or isinstance(self.tf, TransferFunctionTexture):
            return
        new_tf = TransferFunctionTexture()
        new_tf.paste(self.tf)
        self.tf = new_tf


# ------------------------------------------------------------------------------------------------ camera
class Orientation(_Enum): pass


class CameraOnASphere:
    """renderer::CameraOnASphere: pitch/yaw/distance -> (B,3,3) reference frame (camera.cpp:480-581)."""
    Xp, Xm, Yp, Ym, Zp, Zm = (Orientation(n, i) for i, n in enumerate(["Xp", "Xm", "Yp", "Ym", "Zp", "Zm"]))

    def __init__(self):
        self.orientation = CameraOnASphere.Ym
        self.center = Parameter(double3(0, 0, 0))
        self.pitchYawDistance = Parameter(double3(0, 0, 1))
        self.fovY_radians = math.radians(45.0)
        self._matrix: Optional[torch.Tensor] = None

    @property
    def fov_y_radians(self): return self.fovY_radians

    @staticmethod
    def reference_frame(pitch, yaw, distance, center=(0, 0, 0), orientation_index=3) -> np.ndarray:
        """eulerToCartesian + computeParameters + updateCameraMatrix (camera.cpp:485-509, 553-581)."""
        from .scenes import camera_reference_frame
        return camera_reference_frame(pitch, yaw, distance, center, orientation_index)

    def set_parameters(self, t: torch.Tensor):
        if t.dim() != 3 or t.shape[1:] != (3, 3):
            raise RuntimeError("camera parameters must be of shape (B,3,3)")
        self._matrix = t.to(device="cuda", dtype=torch.float32).contiguous()

    def generate_rays(self, width: int, height: int, double_precision: bool = False):
        """ICamera::generateRays (camera.cpp:204-209 -> CameraGenerateRayKernel): (ray_start, ray_dir), both (B,H,W,3)."""
        if double_precision:
            raise RuntimeError("double precision rays are not part of the B200 path")
        cam = self.get_parameters()
        B = cam.shape[0]
        start = torch.empty((B, height, width, 3), dtype=torch.float32, device="cuda")
        dirs = torch.empty_like(start)
        check(_lib.lib().fvsrn_generate_rays(cam.data_ptr(), B, float(self.fovY_radians), width, height, start.data_ptr(), dirs.data_ptr(), _stream_ptr()))
        return start, dirs

    def get_parameters(self) -> torch.Tensor:
        if self._matrix is not None:
            return self._matrix
        pyd, c = self.pitchYawDistance.value, self.center.value
        m = CameraOnASphere.reference_frame(pyd.x, pyd.y, pyd.z, tuple(c), self.orientation.index)
        return torch.from_numpy(m).to("cuda")


# ------------------------------------------------------------------------------------------------ image evaluator
class ChannelMode(_Enum): pass


class IImageEvaluator:
    """IImageEvaluator::ChannelMode (iimage_evaluator.h:19-26, pybind iimage_evaluator.cpp:335-340)."""
    ChannelMode = ChannelMode
    Mask, Normal, Depth, Color = (ChannelMode(n, i) for i, n in enumerate(["Mask", "Normal", "Depth", "Color"]))


OUTPUT_FORMATS = {"raw8_f32": 0, "rgba_f32": 1, "rgba8": 2, "rgba_f16": 3}


class ImageEvaluatorSimple:
    """renderer::ImageEvaluatorSimple::render (image_evaluator_simple.cpp:198-361) -> torch (B,8,H,W) float32."""

    def __init__(self):
        self.camera = CameraOnASphere()
        self.ray_evaluator = RayEvaluationSteppingDvr()
        self.volume: Optional[VolumeInterpolationNetwork] = None
        self.samples_per_pixel = 1
        self.selected_channel = IImageEvaluator.Color
        # tonemapping state, image_evaluator_simple.cpp:21-25, 363-369, pybind :447-451
        self.use_tonemapping = False
        self.fix_max_exposure = False
        self.fixed_max_exposure = 1.0
        self.tonemapping_shoulder = 1.0
        self._last_frame: Optional[torch.Tensor] = None
        self._refining_counter = 0
        self.spp_log2 = 2                    # samplesPerIterationLog2_ (image_evaluator_simple.h): only iterative evaluators use it
        # B200 additions (all optional): screen-tile sharding, per-view times, statistics
        self.tile = None                     # (x0, y0, w, h)
        self.view_times = None               # list of absolute times, one per camera batch entry
        self.collect_statistics = False
        self.last_step_counts: Optional[torch.Tensor] = None
        self.last_counters: Optional[torch.Tensor] = None

    def _volume(self):
        v = self.volume if self.volume is not None else self.ray_evaluator.volume
        if v is None:
            raise RuntimeError("No volume specified")
        return v

    def _params(self, width, height, cam, tf, out_strides):
        re = self.ray_evaluator
        p = _lib.DvrParams()
        p.width, p.height, p.batch = width, height, cam.shape[0]
        if self.tile is not None:
            p.tile_x0, p.tile_y0, p.tile_w, p.tile_h = self.tile
        p.camera_matrix_dev = cam.data_ptr()
        p.fov_y_radians = self.camera.fovY_radians
        p.stepsize, p.alpha_early_out = re.stepsize, re.alpha_early_out
        p.density_min, p.density_max = re.min_density, re.max_density
        p.early_out, p.blend_mode = int(bool(re.early_out)), re.blending.blendMode.index
        if tf is not None:
            p.tf_kind, p.tf_dev = re.tf.kind, tf.data_ptr()
            p.tf_resolution, p.tf_batch = (tf.shape[1] if re.tf.kind != 2 else 0), tf.shape[0]
        p.samples_per_pixel = self.samples_per_pixel
        p.out_stride_b, p.out_stride_c, p.out_stride_y, p.out_stride_x = out_strides
        if re.brdf is not None:
            re.brdf._fill(p, cam)
        return p

    def _render_grid(self, vol: "VolumeInterpolationGrid", width, height, out, fmt=0):
        v = vol._require(); v._ensure_prepared(vol.feature)
        cam = self.camera.get_parameters()
        tf = self.ray_evaluator.tf._tensor()
        B = cam.shape[0]
        shape, dtype = {0: ((B, 8, height, width), torch.float32), 1: ((B, 4, height, width), torch.float32), 2: ((B, height, width), torch.int32),
                        3: ((B, height, width, 4), torch.float16)}[fmt]
        if out is None or tuple(out.shape) != shape or out.dtype != dtype or not out.is_cuda:
            out = (torch.empty if self.tile is None else torch.zeros)(shape, dtype=dtype, device="cuda")
        strides = out.stride() if fmt <= 1 else (out.stride(0) // (4 if fmt == 3 else 1), 0, out.stride(1) // (4 if fmt == 3 else 1), 1)
        p = self._params(width, height, cam, tf, strides)
        p.output_format, p.channel_mode = fmt, self._channel_index()
        p.tonemap, p.max_exposure = int(bool(self.use_tonemapping)), float(self.get_exposure_for_tonemapping() if self.use_tonemapping else 1.0)
        if self.collect_statistics:
            self.last_step_counts = torch.zeros((B, height, width), dtype=torch.int32, device="cuda")
            self.last_counters = torch.zeros(2, dtype=torch.int64, device="cuda")
            p.step_count_dev = self.last_step_counts.data_ptr(); p.counters_dev = self.last_counters.data_ptr()
        check(_lib.lib().fvsrn_render_dvr_grid(v._h, C.byref(p), out.data_ptr(), _stream_ptr()))
        self._refining_counter = 0
        if fmt == 0:
            self._last_frame = out
        return out

    def render(self, width: int, height: int, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        vol = self._volume()
        if isinstance(vol, VolumeInterpolationGrid):
            return self._render_grid(vol, width, height, out)
        net = vol._require()
        net._ensure_prepared()
        vol._apply_gradient_mode(net)
        cam = self.camera.get_parameters()
        is_rgbo = net._info().channels_out == 4
        tf = None if is_rgbo else self.ray_evaluator.tf._tensor()
        B = cam.shape[0]
        if out is None or tuple(out.shape) != (B, 8, height, width) or out.dtype != torch.float32 or not out.is_cuda:
            out = torch.empty((B, 8, height, width), dtype=torch.float32, device="cuda") if self.tile is None \
                else torch.zeros((B, 8, height, width), dtype=torch.float32, device="cuda")
        p = self._params(width, height, cam, tf, out.stride())
        keep = []
        if self.view_times is not None:
            vt = (C.c_float * B)(*[float(t) for t in self.view_times]); keep.append(vt)
            p.view_times_host = C.cast(vt, C.c_void_p)
        if self.collect_statistics:
            self.last_step_counts = torch.zeros((B, height, width), dtype=torch.int32, device="cuda")
            self.last_counters = torch.zeros(2, dtype=torch.int64, device="cuda")
            p.step_count_dev = self.last_step_counts.data_ptr(); p.counters_dev = self.last_counters.data_ptr()
        check(_lib.lib().fvsrn_render_dvr(net._h, C.byref(p), out.data_ptr(), _stream_ptr()))
        self._last_frame = out
        self._refining_counter = 0
        return out

    def refine(self, width: int, height: int, previous: torch.Tensor) -> torch.Tensor:
        """IImageEvaluator.refine (iimage_evaluator.cpp:349-353 -> ImageEvaluatorSimple::render with refine = true,
        image_evaluator_simple.cpp:288-356): render again and fold the frame into the running mean `previous` holds,
        `previous + (frame - previous) / n` with n counted since the last render(). DVR draws no random numbers
        (is_iterative_refining() is False), so every frame equals the first; the bookkeeping is the reference's."""
        if previous is None or tuple(previous.shape) != (self.camera.get_parameters().shape[0], 8, height, width):
            raise RuntimeError("refine: `previous` must be the (B,8,H,W) tensor of the preceding render / refine call")
        counter = self._refining_counter
        frame = self.render(width, height)      # resets the counter, as a non-refining render does
        self._refining_counter = counter + 1
        out = previous + (frame - previous) * (1.0 / self._refining_counter)
        self._last_frame = out
        return out

    def is_iterative_refining(self) -> bool:
        return False                            # RayEvaluationSteppingDvr, ray_evaluation.h: only the Monte-Carlo evaluator refines

    def compute_batch_count(self) -> int:
        """IImageEvaluator::computeBatchCount (iimage_evaluator.cpp:138-165): the common batch size of camera and TF."""
        b = self.camera.get_parameters().shape[0]
        tf = self.ray_evaluator.tf
        if tf is not None and tf.kind != 2 and getattr(tf, "tensor", None) is not None and tf.tensor.value is not None:
            tb = tf.tensor.value.shape[0]
            if tb > 1 and b > 1 and tb != b:
                raise RuntimeError(f"Batch counts don't match. Module camera has {b} batches, but the current module tf has {tb} batches")
            b = max(b, tb)
        return b

    # -- colour extraction (IImageEvaluator::ExtractColor, iimage_evaluator.cpp:26-135)
    @property
    def last_max_exposure(self) -> float:
        """max over rgb of the last rendered frame (image_evaluator_simple.cpp:358; computed on demand here instead of a
        device synchronisation in every render call)."""
        if self._last_frame is None:
            return 0.0
        return float(self._last_frame[:, 0:3].max())

    def get_exposure_for_tonemapping(self) -> float:
        e = self.fixed_max_exposure if self.fix_max_exposure else self.last_max_exposure
        return max(0.001, e * self.tonemapping_shoulder)

    def _channel_index(self):
        c = self.selected_channel
        if isinstance(c, str):      # "ChannelColor" etc. (json names)
            c = {"ChannelMask": 0, "ChannelNormal": 1, "ChannelDepth": 2, "ChannelColor": 3}[c]
        return c.index if isinstance(c, _Enum) else int(c)

    @staticmethod
    def Extract_color(img: torch.Tensor, use_tonemapping: bool, max_exposure: float, channel, output_format="rgba_f32") -> torch.Tensor:
        """(B,8,H,W) -> (B,4,H,W) fp32 (or packed RGBA8 (B,H,W) int32 / (B,H,W,4) fp16), ImageEvaluatorSimple::extractColorTorch."""
        if img.dim() != 4 or img.shape[1] != 8:
            raise RuntimeError("expected a (B,8,H,W) tensor")
        if not img.is_cuda:
            raise RuntimeError("raw input must reside on the GPU")
        img = img.to(torch.float32).contiguous()
        B, _, H, W = img.shape
        fmt = OUTPUT_FORMATS[output_format]
        if fmt == 1: out = torch.empty((B, 4, H, W), dtype=torch.float32, device=img.device)
        elif fmt == 2: out = torch.empty((B, H, W), dtype=torch.int32, device=img.device)
        elif fmt == 3: out = torch.empty((B, H, W, 4), dtype=torch.float16, device=img.device)
        else: raise RuntimeError("Extract_color writes rgba_f32, rgba8 or rgba_f16")
        ch = channel.index if isinstance(channel, _Enum) else int(channel)
        check(_lib.lib().fvsrn_extract_color_ex(img.data_ptr(), out.data_ptr(), B, H, W, ch, int(bool(use_tonemapping)),
                                                float(max_exposure), fmt, _stream_ptr()))
        return out

    def extract_color(self, img: torch.Tensor, output_format="rgba_f32") -> torch.Tensor:
        """Extracts the selected channel and performs tonemapping (pybind image_evaluator_simple.cpp:463-473)."""
        return ImageEvaluatorSimple.Extract_color(img, self.use_tonemapping, self.get_exposure_for_tonemapping() if self.use_tonemapping else 1.0,
                                                  self._channel_index(), output_format)

    def render_color(self, width: int, height: int, output_format="rgba_f32", out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """B200 addition: render + extract_color in ONE kernel - the selected channel (colour / mask / normal; tonemapping
        with fix_max_exposure) leaves the DVR kernel as (B,4,H,W) fp32, packed RGBA8 (B,H,W) int32 or (B,H,W,4) fp16 instead of
        the 32 B/pixel raw tensor."""
        fmt = OUTPUT_FORMATS[output_format]
        if fmt == 0:
            return self.render(width, height, out)
        if self.use_tonemapping and not self.fix_max_exposure:
            raise RuntimeError("fused tonemapping needs fix_max_exposure (the frame-wide maximum is not known inside the frame)")
        if isinstance(self._volume(), VolumeInterpolationGrid):
            return self._render_grid(self._volume(), width, height, out, fmt)
        vol = self._volume(); net = vol._require(); net._ensure_prepared()
        vol._apply_gradient_mode(net)
        cam = self.camera.get_parameters()
        is_rgbo = net._info().channels_out == 4
        tf = None if is_rgbo else self.ray_evaluator.tf._tensor()
        B = cam.shape[0]
        shape, dtype = {1: ((B, 4, height, width), torch.float32), 2: ((B, height, width), torch.int32), 3: ((B, height, width, 4), torch.float16)}[fmt]
        if out is None or tuple(out.shape) != shape or out.dtype != dtype or not out.is_cuda:
            out = (torch.empty if self.tile is None else torch.zeros)(shape, dtype=dtype, device="cuda")
        if fmt == 1: strides = out.stride()
        else: strides = (out.stride(0) // (4 if fmt == 3 else 1), 0, out.stride(1) // (4 if fmt == 3 else 1), 1)
        p = self._params(width, height, cam, tf, strides)
        p.output_format, p.channel_mode = fmt, self._channel_index()
        p.tonemap, p.max_exposure = int(bool(self.use_tonemapping)), float(self.get_exposure_for_tonemapping() if self.use_tonemapping else 1.0)
        keep = []
        if self.view_times is not None:
            vt = (C.c_float * B)(*[float(t) for t in self.view_times]); keep.append(vt)
            p.view_times_host = C.cast(vt, C.c_void_p)
        check(_lib.lib().fvsrn_render_dvr(net._h, C.byref(p), out.data_ptr(), _stream_ptr()))
        return out


# ------------------------------------------------------------------------------------------------ settings files
def load_from_json(filename: str) -> ImageEvaluatorSimple:
    """pyrenderer.load_from_json (renderer/module_registry.cpp:75-99): a scene settings file -> a new ImageEvaluatorSimple with the
    camera, DVR ray evaluator, transfer function, BRDF and volume the file selects, each loaded as that module's `load` does
    (fv-srn_b200/settings.py cites them). A `Grid` volume loads its .cvol (path relative to the settings file); an `SRN` volume comes
    back without networks, as in the reference (set_network / load_network afterwards)."""
    from . import settings
    plan = settings.read_settings(filename)
    ie = ImageEvaluatorSimple()
    e = plan["evaluator"]
    ie.selected_channel = (IImageEvaluator.Mask, IImageEvaluator.Normal, IImageEvaluator.Depth, IImageEvaluator.Color)[e["channel"]]
    ie.spp_log2 = e["samples_per_iteration_log2"]
    ie.use_tonemapping, ie.tonemapping_shoulder = e["use_tonemapping"], e["tonemapping_shoulder"]
    ie.fix_max_exposure, ie.fixed_max_exposure = e["fix_max_exposure"], e["fixed_max_exposure"]
    c = plan["camera"]
    cam = ie.camera
    cam.orientation = (CameraOnASphere.Xp, CameraOnASphere.Xm, CameraOnASphere.Yp, CameraOnASphere.Ym, CameraOnASphere.Zp, CameraOnASphere.Zm)[c["orientation"]]
    cam.center = Parameter(double3(*c["center"]))
    cam.pitchYawDistance = Parameter(double3(c["pitch"], c["yaw"], c["distance"]))
    cam.fovY_radians = c["fov_y"]
    cam.zoom_value = c["zoom"]
    re, d = ie.ray_evaluator, plan["dvr"]
    re.stepsize, re.min_density, re.max_density, re.early_out = d["stepsize"], d["min_density"], d["max_density"], d["early_out"]
    t = plan["tf"]
    if t is None:
        re.tf = None
    elif t["kind"] == "Piecewise":
        re.tf = TransferFunctionPiecewiseLinear()
        re.tf.load_points(t["color_points"], t["opacity_points"], t["absorption_scaling"])
    elif t["kind"] == "Texture":
        re.tf = TransferFunctionTexture()
        re.tf.preintegration_mode = t["preintegration_mode"]
        re.tf.load_points(t["color_points"], t["plot"], t["absorption_scaling"])
    else:
        re.tf = TransferFunctionIdentity(t["absorption_scaling"], t["emission_scaling"])
    b = plan["brdf"]
    if b is None:
        re.brdf = None
    else:
        br = re.brdf = BRDFLambert()
        br.enable_magnitude_scaling, br.magnitude_scaling = b["enable_magnitude_scaling"], b["magnitude_scaling"]
        br.enable_phong, br.ambient, br.specular = b["enable_phong"], b["ambient"], b["specular"]
        br.magnitude_center, br.magnitude_radius, br.specular_exponent = b["magnitude_center"], b["magnitude_radius"], b["specular_exponent"]
        br.light_follows_camera = b["light_follows_camera"]
        br.light_type = (BRDFLambert.Point, BRDFLambert.Directional)[b["light_type"]]
        br.light_position, br.light_direction = double3(*b["light_position"]), double3(*b["light_direction"])
    v = plan["volume"]
    if v is None:
        ie.volume = None
    elif v["kind"] == "Grid":
        g = VolumeInterpolationGrid()
        g.feature = v["density_feature"]
        if v["path"] is not None:
            if v["interpolation"] != "TRILINEAR":
                raise RuntimeError(f"grid interpolation {v['interpolation']} is not part of the B200 path (TRILINEAR is)")
            if not os.path.isfile(v["path"]):       # the reference prints this and leaves the source EMPTY; rendering then fails with "No volume specified"
                import sys
                print(f"Unable to load volume, is the file valid? {v['path']}", file=sys.stderr)
            elif v["path"].endswith(".json"):
                raise RuntimeError("ensemble volume lists are not part of the B200 path")
            else:
                g.setSource(Volume(v["path"]), v["mipmap_level"])
        ie.volume = g
    else:
        n = VolumeInterpolationNetwork()
        n.gradient_mode = (VolumeInterpolationNetwork.OFF_OR_DIRECT, VolumeInterpolationNetwork.FINITE_DIFFERENCES,
                           VolumeInterpolationNetwork.ADJOINT_METHOD)[v["gradient_mode"]]
        n.finite_differences_stepsize = v["finite_differences_stepsize"]
        ie.volume = n
    return ie
