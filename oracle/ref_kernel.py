"""
TEST INFRASTRUCTURE — runs the REFERENCE's own fused DVR kernel on the GPU box.

oracle/build_ref_ptx.py assembled the reference's ImageEvaluatorSimpleKernel (renderer/*.cuh, unmodified) into PTX in the
build container; this module loads that PTX through the CUDA driver API (the reference does the same with
cuModuleLoadDataEx, renderer/kernel_loader.cpp:496-556), fills the reference's __constant__ blocks the way its host code
does (SceneNetwork::fillConstantMemory, renderer/volume_interpolation_network.cpp:1236-1417 — every member padded to
32 bytes; camera / DVR / TF blocks per SURVEY App. B), uploads the latent grid as C/4 float4 3-D textures (clamp, linear,
normalised — LatentGrid::copyGridToGPU, :470-504) and launches with the reference's launch shape
(image_evaluator_simple.cpp:300-345: 1-D grid of BLOCK_SIZE threads, capped by occupancy).
Used only by tests/ (parity comparand) and bench.py (GPU baseline). Nothing here is product code.
"""
import ctypes as C
import json
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(HERE, "_ref")


def available():
    return os.path.exists(os.path.join(REF_DIR, "ref_kernels.json"))


def _ck(res):
    from cuda.bindings import driver
    err = res[0]
    if err != driver.CUresult.CUDA_SUCCESS:
        raise RuntimeError(f"CUDA driver error {err}")
    return res[1] if len(res) == 2 else (res[1:] if len(res) > 2 else None)


def _pad32(buf: bytearray):
    while len(buf) % 32:
        buf.append(0)


class _Accessor3(C.Structure):   # kernel::Tensor3Read<float>: {T* data; int32 sizes[3]; int32 strides[3]} (renderer_tensor.cuh:123-183)
    _fields_ = [("data", C.c_uint64), ("sizes", C.c_int32 * 3), ("strides", C.c_int32 * 3)]


class _Accessor4(C.Structure):   # kernel::Tensor4RW<float>
    _fields_ = [("data", C.c_uint64), ("sizes", C.c_int32 * 4), ("strides", C.c_int32 * 4)]


class _CameraParams(C.Structure):  # kernel::CameraReferenceFrameParameters (renderer_camera.cuh:12-25)
    _fields_ = [("fovY", C.c_float), ("aspect", C.c_float), ("matrix", _Accessor3)]


class _Dim3(C.Structure):
    _fields_ = [("x", C.c_uint32), ("y", C.c_uint32), ("z", C.c_uint32)]


def _read3(t, broadcast=True):
    a = _Accessor3()
    a.data = t.data_ptr()
    for i in range(3):
        a.sizes[i] = t.shape[i]
        a.strides[i] = 0 if (broadcast and t.shape[i] == 1) else t.stride(i)
    return a


class ReferenceDvrKernel:
    """One JIT-loaded reference kernel bound to one network (density network, Fourier features, one 16-channel latent
    grid group, no direction/time *input*). Covers the paper-default shape, the 64x6 network of C4,
    time-interpolated keyframe grids (C5: cLatentGridB != A, cLatentGridInterpolation = t), byte-encoded grids and the
    texture transfer function (tensor form)."""

    OUTPUT_MODES = ["density", "density:direct", "rgbo", "rgbo:direct", "densitygrad", "densitygrad:direct", "densitygrad:cubic",
                    "densitycurvature", "densitycurvature:direct"]

    def __init__(self, net, variant=None, patched_steps=False, time=0.0, tf="piecewise", fd_stepsize=0.01):
        import torch
        from cuda.bindings import driver
        self.torch, self.drv = torch, driver
        meta = json.load(open(os.path.join(REF_DIR, "ref_kernels.json")))
        act = net.layers[0].activation
        H = net.input_channels_out()
        grids = net.time_grids or net.ensemble_grids
        enc = grids[0].encoding
        if variant is None:
            variant = f"h{H}x{len(net.layers)}_{act.lower()}" + {0: "", 1: "_bytelinear", 2: "_bytegauss"}[enc] \
                      + {"piecewise": "", "texture": "_textf", "identity": "_identitytf"}[tf] + ("_steps" if patched_steps else "")
        if variant not in meta:
            raise RuntimeError(f"no prebuilt reference PTX for {variant} (have {sorted(meta)})")
        m = meta[variant]
        assert m["hidden16"] * 16 == H and m["num_hidden"] == len(net.layers) - 2 and m["nf"] == net.num_fourier
        assert net.latent_channels() == 16 * m.get("grid16", 1) and int(net.has_time) == m.get("pass_time", 0)
        assert (0 if not net.has_direction else (2 if net.direction_in_fourier else 1)) == m.get("use_direction", 0)
        assert net.output_mode == self.OUTPUT_MODES[m["output_mode"]], (net.output_mode, m["output_mode"])
        self.fd_stepsize = float(fd_stepsize)
        assert m.get("encoding", 0) == enc and m.get("tf", "piecewise") == tf
        self.meta, self.block, self.tf_kind = m, m["block_size"], tf
        torch.cuda.init(); torch.zeros(1, device="cuda")         # primary context current
        ptx = open(os.path.join(REF_DIR, m["ptx"]), "rb").read() + b"\0"
        self.module = _ck(driver.cuModuleLoadData(ptx))          # JIT compute_75 PTX -> sm_100 SASS
        low = m["lowered"]
        self.g_vol = _ck(driver.cuModuleGetGlobal(self.module, low["&volumeInterpolationTensorcoresParameters"].encode()))
        self._keep = []
        if m.get("eval"):       # the batch-query build (ReferenceEvalKernel): no camera / marcher / TF constants
            self.func = _ck(driver.cuModuleGetFunction(self.module, low["EvaluateNoBatches"].encode()))
            self._upload_network(net, float(time), unit_box=True)
            self._occupancy()
            return
        self.func = _ck(driver.cuModuleGetFunction(self.module, low["ImageEvaluatorSimpleKernel"].encode()))
        self.g_cam = _ck(driver.cuModuleGetGlobal(self.module, low["&cameraReferenceFrameParameters"].encode()))
        self.g_dvr = _ck(driver.cuModuleGetGlobal(self.module, low["&rayEvaluationSteppingDvrParameters"].encode()))
        tf_sym = {"texture": "&transferFunctionTextureParameters", "identity": "&transferFunctionIdentityParameters"}.get(tf, "&transferFunctionPiecewiseParameters")
        self.g_tf = _ck(driver.cuModuleGetGlobal(self.module, low[tf_sym].encode()))
        # BRDFLambertParameters (scalar form, renderer_brdf_lambert.cuh:30-38); absent from the module when the BRDF is the pass-through
        self.g_brdf = None
        if m.get("brdf") and "&brdfLambertParameters" in low:
            self.g_brdf = _ck(driver.cuModuleGetGlobal(self.module, low["&brdfLambertParameters"].encode()))
        self._upload_network(net, float(time))
        self._occupancy()

    def _occupancy(self):
        sms = self.torch.cuda.get_device_properties(0).multi_processor_count
        bps = _ck(self.drv.cuOccupancyMaxActiveBlocksPerMultiprocessor(self.func, self.block, 0))
        self.max_grid = max(1, bps) * sms
        self.blocks_per_sm = bps

    # ---- latent grid -> C/4 4-channel 3-D textures (LatentGrid::GPUArray, volume_interpolation_network.cpp:470-504)
    def _texture(self, vol4):
        """vol4: (Z,Y,X,4) float32 (element-type read) or uint8 (normalised-float read)."""
        d = self.drv
        Z, Y, X, _ = vol4.shape
        is_float = vol4.dtype != np.uint8
        bpt = 4 if is_float else 1
        desc = d.CUDA_ARRAY3D_DESCRIPTOR()
        desc.Width, desc.Height, desc.Depth = X, Y, Z
        desc.Format = d.CUarray_format.CU_AD_FORMAT_FLOAT if is_float else d.CUarray_format.CU_AD_FORMAT_UNSIGNED_INT8
        desc.NumChannels = 4
        desc.Flags = 0
        arr = _ck(d.cuArray3DCreate(desc))
        host = np.ascontiguousarray(vol4, np.float32 if is_float else np.uint8)
        cp = d.CUDA_MEMCPY3D()
        cp.srcMemoryType = d.CUmemorytype.CU_MEMORYTYPE_HOST
        cp.srcHost = host.ctypes.data
        cp.srcPitch = X * 4 * bpt
        cp.srcHeight = Y
        cp.dstMemoryType = d.CUmemorytype.CU_MEMORYTYPE_ARRAY
        cp.dstArray = arr
        cp.WidthInBytes = X * 4 * bpt
        cp.Height = Y
        cp.Depth = Z
        _ck(d.cuMemcpy3D(cp))
        res = d.CUDA_RESOURCE_DESC()
        res.resType = d.CUresourcetype.CU_RESOURCE_TYPE_ARRAY
        res.res.array.hArray = arr
        td = d.CUDA_TEXTURE_DESC()
        td.addressMode = [d.CUaddress_mode.CU_TR_ADDRESS_MODE_CLAMP] * 3
        td.filterMode = d.CUfilter_mode.CU_TR_FILTER_MODE_LINEAR
        td.flags = d.CU_TRSF_NORMALIZED_COORDINATES      # (no CU_TRSF_READ_AS_INTEGER: 8-bit texels read as normalised floats)
        tex = _ck(d.cuTexObjectCreate(res, td, None))
        self._keep.append((arr, host))
        return int(tex)

    def _grid_textures(self, grid):
        if grid.encoding == 0:
            # the same fp16-rounded texels every other implementation sees, stored as fp32 like the reference
            vals = grid.values.astype(np.float32).astype(np.float16).astype(np.float32)     # (C,Z,Y,X)
        else:
            vals = grid.raw.astype(np.uint8)
        return [self._texture(np.ascontiguousarray(np.transpose(vals[4 * i:4 * i + 4], (1, 2, 3, 0)))) for i in range(vals.shape[0] // 4)]

    def _upload_network(self, net, time, unit_box=False):
        H, nf = net.input_channels_out(), net.num_fourier
        grids = net.time_grids or net.ensemble_grids
        # keyframe selection as SceneNetwork::setTimeAndEnsemble / fillConstantMemory (:1290-1352): `time` is the keyframe index
        T = len(grids)
        lo = min(int(time), T - 1); hi = min(lo + 1, T - 1)
        gA, gB = grids[lo], grids[hi]
        texA = self._grid_textures(gA)
        texB = texA if hi == lo else self._grid_textures(gB)
        b = bytearray()

        def add(x):
            _pad32(b); b.extend(x)

        add(np.ascontiguousarray(net.fourier.T, np.float16).tobytes())                  # cWeightsFourier [cin][nf]
        add(np.array(texA, np.uint64).tobytes())                                        # cLatentGridA
        add(np.array(texB, np.uint64).tobytes())                                        # cLatentGridB
        if gA.encoding != 0:                                                            # HAS_LATENT_GRID_OFFSET_SCALE
            add(np.asarray(gA.offset_or_mean, np.float32).tobytes()); add(np.asarray(gB.offset_or_mean, np.float32).tobytes())
            add(np.asarray(gA.scale_or_std, np.float32).tobytes()); add(np.asarray(gB.scale_or_std, np.float32).tobytes())
        add(np.full(net.latent_channels(), time, np.float32).tobytes())                 # cLatentGridInterpolation (kernel takes frac; [0] = the time input)
        add(np.ascontiguousarray(net.layers[0].weights, np.float16).tobytes())          # cWeightsLatentGrid [H][H+C]
        add(np.ascontiguousarray(net.layers[0].bias, np.float16).tobytes())
        if len(net.layers) > 2:
            add(np.ascontiguousarray(np.stack([l.weights for l in net.layers[1:-1]]), np.float16).tobytes())
            add(np.ascontiguousarray(np.stack([l.bias for l in net.layers[1:-1]]), np.float16).tobytes())
        else:       # NUM_HIDDEN_LAYERS 0: the device struct keeps one element (MIN_ONE, renderer_volume_tensorcores.cuh:223-224)
            add(np.zeros(1, np.float16).tobytes()); add(np.zeros(1, np.float16).tobytes())
        add(np.ascontiguousarray(np.asarray(net.layers[-1].weights).T, np.float16).tobytes())   # cWeightsLast [H] / half2 [H][cout/2]
        add(np.ascontiguousarray(net.layers[-1].bias, np.float16).tobytes())
        # IVolumeInterpolation::evaluate forces the box to the unit cube (volume_interpolation.cpp:45-48)
        add(np.array((0, 0, 0) if unit_box else net.box_min, np.float32).tobytes())
        add(np.array((1, 1, 1) if unit_box else net.box_size, np.float32).tobytes())
        add(np.array([net.layers[0].activation_param], np.float32).tobytes())
        if self.meta.get("gradient_mode", 0) == 2:
            # VolumeInterpolationNetwork::fillConstantMemory (volume_interpolation_network.cpp:1809-1812): in adjoint mode both slots
            # get 1 / (float(gridResolution) * adjointLatentGridCentralDifferencesStepsizeScale_), scale 4 (:1424)
            step = np.float32(1.0) / (np.float32(gA.values.shape[-1]) * np.float32(4.0))
            add(np.array([step], np.float32).tobytes()); add(np.array([step], np.float32).tobytes())
        else:
            add(np.array([self.fd_stepsize], np.float32).tobytes())                     # finiteDifferencesStepsize (GRADIENT_MODE 1)
            add(np.array([0.01], np.float32).tobytes())                                 # latentGridDifferencesStepSize (unused)
        _pad32(b)
        dptr, size = self.g_vol
        if H == 32 and nf == 14 and gA.encoding == 0 and net.layers[-1].cout == 1 and net.latent_channels() == 16:
            assert len(b) == 7840, len(b)                                               # sizeof for C3
        assert len(b) == int(size), (len(b), int(size))
        _ck(self.drv.cuMemcpyHtoD(dptr, bytes(b), len(b)))

    def render(self, width, height, camera, fov_y, stepsize, tf, alpha_early_out=1 - 1e-5, density_min=0.0, density_max=1.0,
               out=None, timing_iters=0, brdf=None):
        """-> (B,8,H,W) float32 CUDA tensor; with timing_iters > 0 also the mean kernel time in ms (CUDA events).
        tf: (1|B,R,5) piecewise or (1|B,R,4) texture tensor, matching the variant this kernel was built with."""
        torch, d = self.torch, self.drv
        cam = torch.as_tensor(np.ascontiguousarray(camera, np.float32)).reshape(-1, 3, 3).cuda()
        if self.tf_kind == "identity":      # TransferFunctionIdentityParameters {scaleAbsorption, scaleEmission} (renderer_tf_identity.cuh:16-26)
            tft = None
            tf_blob = np.asarray(tf, np.float32).reshape(-1)[:2].tobytes()
        else:
            tft = torch.as_tensor(np.ascontiguousarray(tf, np.float32)).cuda()
            tft = tft.reshape(-1, tft.shape[-2], tft.shape[-1])
            assert tft.shape[-1] == (4 if self.tf_kind == "texture" else 5)
        B = cam.shape[0]
        if (width * height * B) % 32:
            raise RuntimeError("the reference requires W*H*B to be a multiple of 32 (image_evaluator_simple.cpp:254-264)")
        cp = _CameraParams(); cp.fovY = fov_y; cp.aspect = float(width) / float(height); cp.matrix = _read3(cam)
        _ck(d.cuMemcpyHtoD(self.g_cam[0], bytes(cp), C.sizeof(cp)))
        _ck(d.cuMemcpyHtoD(self.g_dvr[0], np.array([stepsize, alpha_early_out, density_min, density_max], np.float32).tobytes(), 16))
        if tft is None:
            _ck(d.cuMemcpyHtoD(self.g_tf[0], tf_blob, 8))
        else:
            ta = _read3(tft)
            _ck(d.cuMemcpyHtoD(self.g_tf[0], bytes(ta), C.sizeof(ta)))
        if self.g_brdf is not None:
            # brdf: dict(magnitude_scaling, ambient, specular, magnitude_center, magnitude_radius, specular_exponent, light) - light =
            # position (point light) or direction (directional), as BRDFLambert::fillConstantMemory packs it (brdf.cpp:417-452)
            blob = np.array([brdf["magnitude_scaling"], brdf["ambient"], brdf["specular"], brdf["magnitude_center"], brdf["magnitude_radius"]],
                            np.float32).tobytes() + np.array([brdf["specular_exponent"]], np.int32).tobytes() + np.asarray(brdf["light"], np.float32).tobytes()
            assert len(blob) == 36 == int(self.g_brdf[1]), (len(blob), int(self.g_brdf[1]))
            _ck(d.cuMemcpyHtoD(self.g_brdf[0], blob, len(blob)))
        if out is None:
            out = torch.empty((B, 8, height, width), dtype=torch.float32, device="cuda")
        oa = _Accessor4(); oa.data = out.data_ptr()
        for i in range(4):
            oa.sizes[i] = out.shape[i]; oa.strides[i] = out.stride(i)
        vs = _Dim3(width, height, B)
        samples, time, bg = C.c_int(1), C.c_uint(0), _Accessor4()
        args = (C.c_void_p * 5)(C.addressof(vs), C.addressof(oa), C.addressof(samples), C.addressof(time), C.addressof(bg))
        grid = min((width * height * B + self.block - 1) // self.block, self.max_grid)
        stream = torch.cuda.current_stream().cuda_stream

        def launch():
            _ck(d.cuLaunchKernel(self.func, grid, 1, 1, self.block, 1, 1, 0, stream, C.addressof(args), 0))

        launch()
        torch.cuda.synchronize()
        if timing_iters <= 0:
            return out
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ms = 0.0
        flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
        for _ in range(timing_iters):
            flush.zero_()
            e0.record(); launch(); e1.record(); torch.cuda.synchronize()
            ms += e0.elapsed_time(e1)
        return out, ms / timing_iters


class _Accessor2(C.Structure):   # kernel::Tensor2RW<float>
    _fields_ = [("data", C.c_uint64), ("sizes", C.c_int32 * 2), ("strides", C.c_int32 * 2)]


class ReferenceEvalKernel(ReferenceDvrKernel):
    """The reference's world-space batch query on the GPU: IVolumeInterpolation::evaluate -> EvaluateNoBatches
    (volume_interpolation.cpp:26-127, renderer_volume_kernels1.cuh:15-71) with the tensor-core SRN, built by
    oracle/build_ref_ptx.py as `<network>_eval`. positions (N,3) in the unit cube -> densities (N,1)."""

    def __init__(self, net, time=0.0):
        act = net.layers[0].activation
        mode = {"density:direct": "", "rgbo": "_rgbo", "rgbo:direct": "_rgbo_direct"}[net.output_mode]
        self.channels = 4 if mode else 1
        super().__init__(net, variant=f"h{net.input_channels_out()}x{len(net.layers)}_{act.lower()}{mode}_eval", time=time)

    def evaluate(self, positions, timing_iters=0):
        torch, d = self.torch, self.drv
        pos = positions if torch.is_tensor(positions) else torch.as_tensor(np.ascontiguousarray(positions, np.float32))
        pos = pos.to("cuda", torch.float32).contiguous()
        n = pos.shape[0]
        out = torch.empty((n, self.channels), dtype=torch.float32, device="cuda")

        def acc(t):
            a = _Accessor2(); a.data = t.data_ptr()
            for i in range(2):
                a.sizes[i] = t.shape[i]; a.strides[i] = t.stride(i)
            return a
        vs, pa, da, oa = _Dim3(n, 1, 1), acc(pos), _Accessor2(), acc(out)
        args = (C.c_void_p * 4)(C.addressof(vs), C.addressof(pa), C.addressof(da), C.addressof(oa))
        grid = min((n + self.block - 1) // self.block, self.max_grid)
        stream = torch.cuda.current_stream().cuda_stream

        def launch():
            _ck(d.cuLaunchKernel(self.func, grid, 1, 1, self.block, 1, 1, 0, stream, C.addressof(args), 0))
        launch(); torch.cuda.synchronize()
        if timing_iters <= 0:
            return out
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
        ms = 0.0
        for _ in range(timing_iters):
            flush.zero_(); e0.record(); launch(); e1.record(); torch.cuda.synchronize(); ms += e0.elapsed_time(e1)
        return out, ms / timing_iters


class ReferenceGridDvrKernel:
    """The reference's DVR kernel over a ground-truth density grid (VolumeInterpolationGrid, texture path, trilinear;
    oracle/build_ref_ptx.py grid_source): same camera / DVR / TF constant blocks, volumeInterpolationGridParameters =
    {tex; int3 resolution-1; float3 boxMin, boxSize, normalStep, normalScale} (renderer_volume_grid.cuh:26-40, host side
    volume_interpolation_grid.cpp:1050-1110). density: (Z,Y,X) float32 or uint8 (read as normalised float)."""

    def __init__(self, density, world_size=(1.0, 1.0, 1.0), patched_steps=False):
        import torch
        from cuda.bindings import driver as d
        self.torch, self.drv = torch, d
        meta = json.load(open(os.path.join(REF_DIR, "ref_kernels.json")))
        m = meta["grid_trilinear_steps" if patched_steps else "grid_trilinear"]
        torch.cuda.init(); torch.zeros(1, device="cuda")
        self.module = _ck(d.cuModuleLoadData(open(os.path.join(REF_DIR, m["ptx"]), "rb").read() + b"\0"))
        low = m["lowered"]
        self.func = _ck(d.cuModuleGetFunction(self.module, low["ImageEvaluatorSimpleKernel"].encode()))
        self.g_cam = _ck(d.cuModuleGetGlobal(self.module, low["&cameraReferenceFrameParameters"].encode()))
        self.g_vol = _ck(d.cuModuleGetGlobal(self.module, low["&volumeInterpolationGridParameters"].encode()))
        self.g_dvr = _ck(d.cuModuleGetGlobal(self.module, low["&rayEvaluationSteppingDvrParameters"].encode()))
        self.g_tf = _ck(d.cuModuleGetGlobal(self.module, low["&transferFunctionPiecewiseParameters"].encode()))
        host = np.ascontiguousarray(density)
        Z, Y, X = host.shape
        is_float = host.dtype != np.uint8
        desc = d.CUDA_ARRAY3D_DESCRIPTOR()
        desc.Width, desc.Height, desc.Depth = X, Y, Z
        desc.Format = d.CUarray_format.CU_AD_FORMAT_FLOAT if is_float else d.CUarray_format.CU_AD_FORMAT_UNSIGNED_INT8
        desc.NumChannels = 1; desc.Flags = 0
        arr = _ck(d.cuArray3DCreate(desc))
        bpt = 4 if is_float else 1
        cp = d.CUDA_MEMCPY3D()
        cp.srcMemoryType = d.CUmemorytype.CU_MEMORYTYPE_HOST; cp.srcHost = host.ctypes.data; cp.srcPitch = X * bpt; cp.srcHeight = Y
        cp.dstMemoryType = d.CUmemorytype.CU_MEMORYTYPE_ARRAY; cp.dstArray = arr
        cp.WidthInBytes = X * bpt; cp.Height = Y; cp.Depth = Z
        _ck(d.cuMemcpy3D(cp))
        res = d.CUDA_RESOURCE_DESC(); res.resType = d.CUresourcetype.CU_RESOURCE_TYPE_ARRAY; res.res.array.hArray = arr
        td = d.CUDA_TEXTURE_DESC()
        td.addressMode = [d.CUaddress_mode.CU_TR_ADDRESS_MODE_CLAMP] * 3
        td.filterMode = d.CUfilter_mode.CU_TR_FILTER_MODE_LINEAR
        td.flags = 0                                        # unnormalised coordinates; integer texels read as normalised floats
        tex = int(_ck(d.cuTexObjectCreate(res, td, None)))
        self._keep = [arr, host]
        ws = np.asarray(world_size, np.float64)
        voxel = ws / np.array([X, Y, Z], np.float64)
        b = np.array([tex], np.uint64).tobytes() + np.array([X - 1, Y - 1, Z - 1], np.int32).tobytes() \
            + (-ws / 2).astype(np.float32).tobytes() + ws.astype(np.float32).tobytes() \
            + np.ones(3, np.float32).tobytes() + (0.5 / voxel).astype(np.float32).tobytes()
        b += b"\0" * (int(self.g_vol[1]) - len(b))
        assert len(b) == int(self.g_vol[1]) and len(b) in (68, 72), len(b)
        _ck(d.cuMemcpyHtoD(self.g_vol[0], b, len(b)))
        self.block = 256
        sms = torch.cuda.get_device_properties(0).multi_processor_count
        self.max_grid = max(1, _ck(d.cuOccupancyMaxActiveBlocksPerMultiprocessor(self.func, self.block, 0))) * sms

    def render(self, width, height, camera, fov_y, stepsize, tf, alpha_early_out=1 - 1e-5, density_min=0.0, density_max=1.0):
        torch, d = self.torch, self.drv
        cam = torch.as_tensor(np.ascontiguousarray(camera, np.float32)).reshape(-1, 3, 3).cuda()
        tft = torch.as_tensor(np.ascontiguousarray(tf, np.float32)).cuda()
        tft = tft.reshape(-1, tft.shape[-2], tft.shape[-1])
        B = cam.shape[0]
        cp = _CameraParams(); cp.fovY = fov_y; cp.aspect = float(width) / float(height); cp.matrix = _read3(cam)
        _ck(d.cuMemcpyHtoD(self.g_cam[0], bytes(cp), C.sizeof(cp)))
        _ck(d.cuMemcpyHtoD(self.g_dvr[0], np.array([stepsize, alpha_early_out, density_min, density_max], np.float32).tobytes(), 16))
        ta = _read3(tft)
        _ck(d.cuMemcpyHtoD(self.g_tf[0], bytes(ta), C.sizeof(ta)))
        out = torch.empty((B, 8, height, width), dtype=torch.float32, device="cuda")
        oa = _Accessor4(); oa.data = out.data_ptr()
        for i in range(4):
            oa.sizes[i] = out.shape[i]; oa.strides[i] = out.stride(i)
        vs = _Dim3(width, height, B)
        samples, time, bg = C.c_int(1), C.c_uint(0), _Accessor4()
        args = (C.c_void_p * 5)(C.addressof(vs), C.addressof(oa), C.addressof(samples), C.addressof(time), C.addressof(bg))
        grid = min((width * height * B + self.block - 1) // self.block, self.max_grid)
        _ck(d.cuLaunchKernel(self.func, grid, 1, 1, self.block, 1, 1, 0, torch.cuda.current_stream().cuda_stream, C.addressof(args), 0))
        torch.cuda.synchronize()
        return out
