"""
Generates the golden fixtures in tests/golden/ by running the REFERENCE's own Python code on the CPU:
  * applications/volnet/network.py  SceneRepresentationNetwork (random init, fixed seeds) -> fp32 outputs
  * applications/volnet/raytracing.py Raytracing._full_trace_forward                    -> fp32 DVR images
The reference imports ``pyrenderer`` (CUDA-only C++ module, not buildable here: SURVEY F7); it is stubbed in
sys.modules and the three calls raytracing.py makes into it are served by CPU restatements that follow
renderer/renderer_camera.cuh:34-52 (+ renderer_camera_kernels.cuh:12-43), renderer/renderer_tf_kernels.cuh:42-70 +
renderer_tf_piecewise.cuh:27-62, and the network's box.  The export to .volnet follows
SceneRepresentationNetwork.export_to_pyrenderer (network.py:798-897).

Runs only in the build container (needs /root/reference); the outputs are committed:
    python tests/golden/make_golden.py
"""
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = os.environ.get("FVSRN_REFERENCE", "/root/reference")
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(REF, "applications"))

import fvsrn_b200  # noqa: E402
from fvsrn_b200 import scenes, volnet  # noqa: E402


# ------------------------------------------------------------------------------------------------ pyrenderer stub
class _Vec:
    def __init__(self, *v): self.x, self.y, self.z = v[0], v[1], v[2]; self.w = v[3] if len(v) > 3 else 0


class _Stub(types.ModuleType):
    class float3(_Vec): pass
    class float4(_Vec): pass
    class double3(_Vec): pass
    class double4(_Vec): pass

    @staticmethod
    def cleanup(): pass

    def __getattr__(self, name):  # any other attribute: a dummy class (only used in type annotations)
        return type(name, (), {})


sys.modules["pyrenderer"] = _Stub("pyrenderer")
from volnet.network import SceneRepresentationNetwork  # noqa: E402  (the reference)
from volnet.raytracing import Raytracing  # noqa: E402  (the reference)


class _FakeInput:
    def __init__(self, T): self._T = T
    def num_timekeyframes(self): return self._T
    def num_ensembles(self): return 1


def make_opt(layers, activation, fouriercount, fourierstd, grid_res, grid_ch=16, outputmode="density:direct", time_dependent=False):
    # as applications/volnet/collect_possible_layers.py:14-46
    return {
        'outputmode': outputmode, 'layers': layers, 'activation': activation, 'fouriercount': fouriercount,
        'fourierstd': fourierstd, 'time_features': grid_ch if time_dependent else 0, 'ensemble_features': 0,
        'volumetric_features_channels': grid_ch, 'volumetric_features_resolution': grid_res,
        'volumetric_features_std': 0.5, 'volumetric_features_time_dependent': time_dependent,
        'use_direction': False, 'disable_direction_in_fourier_features': False, 'fourier_position_direction_split': -1,
        'use_time_direct': False, 'num_time_fourier': False, 'meta_network': None, 'meta_activation': None,
        'meta_pretrain': None,
    }


def export(net: SceneRepresentationNetwork, opt, box_min, box_size) -> volnet.SceneNetwork:
    """network.py:798-897 with our volnet.SceneNetwork in place of pyrenderer.SceneNetwork."""
    n = volnet.SceneNetwork()
    B = net._input_parametrization.get_fourier_feature_matrix()
    n.has_direction = net.use_direction()
    n.set_fourier_matrix(B.detach().numpy(), net._input_parametrization.is_premultiplied())
    n.has_time = net._input_parametrization.has_time()
    n.output_mode = opt['outputmode']
    if net._volumetric_features_time_dependent:
        grids = net._volumetric_latent_space_time.detach().numpy()
        n.has_latent_grid = True
        n.time_min, n.time_num, n.time_step, n.ensemble_min, n.ensemble_num = 0, grids.shape[0], 1, 0, 0
        n.time_grids = [volnet.LatentGrid(volnet.GRID_FLOAT, np.ascontiguousarray(g)) for g in grids]
    elif net._has_volumetric_features:
        n.has_latent_grid = True
        n.time_min, n.time_num, n.time_step, n.ensemble_min, n.ensemble_num = 0, 1, 1, 0, 0
        n.time_grids = [volnet.LatentGrid(volnet.GRID_FLOAT, np.ascontiguousarray(net._volumetric_latent_space.detach().numpy()[0]))]
    act = opt['activation'].split(':')
    param = float(act[1]) if len(act) >= 2 else 1.0
    sizes = list(map(int, opt['layers'].split(':')))
    for i in range(len(sizes)):
        l = getattr(net._hidden_layers, 'linear%d' % i)
        n.add_layer(l.weight.detach().numpy(), l.bias.detach().numpy(), act[0], param)
    l = getattr(net._hidden_layers, 'linear%d' % len(sizes))
    n.add_layer(l.weight.detach().numpy(), l.bias.detach().numpy(), "None", 1.0)
    n.box_min, n.box_size = tuple(box_min), tuple(box_size)
    assert n.valid()
    return n


def encode_grid(values: np.ndarray, encoding: int) -> volnet.LatentGrid:
    """LatentGrid::initEncodingByteLinear / initEncodingByteGaussian (volume_interpolation_network.cpp:336-431) in NumPy:
    (C,Z,Y,X) float32 -> stored bytes + per-channel constants, and the values they DECODE to per voxel
    (renderer_volume_tensorcores.cuh:341-383)."""
    from scipy.special import erf, erfinv
    v = values.astype(np.float32)
    C_ = v.shape[0]
    raw = np.empty(v.shape, np.uint8); off = np.empty(C_, np.float32); sc = np.empty(C_, np.float32)
    for c in range(C_):
        if encoding == volnet.GRID_BYTE_LINEAR:
            mn, mx = np.float32(v[c].min()), np.float32(v[c].max())
            off[c], sc[c] = mn, mx - mn
            x01 = (v[c] - mn) * (np.float32(1.0) / max(np.float32(1e-5), mx - mn))
        else:
            mean, std = np.float32(v[c].mean()), np.float32(v[c].std(ddof=1))          # torch::std_mean: unbiased
            off[c], sc[c] = mean, std
            vhat = (v[c] - mean) * (np.float32(1.0) / max(np.float32(1e-5), std))
            x01 = (np.float32(0.5) * (1 + erf(vhat * np.float32(0.7071067811865475)))).astype(np.float32)
        raw[c] = np.clip(np.floor(np.float32(255) * x01 + np.float32(0.5)), 0, 255).astype(np.uint8)          # roundf
    u = raw.astype(np.float32) / np.float32(255)
    if encoding == volnet.GRID_BYTE_LINEAR:
        dec = off[:, None, None, None] + u * sc[:, None, None, None]
    else:
        t = (np.sqrt(2.0) * erfinv((np.float32(2.0) - np.float32(1e-4)) * (u - np.float32(0.5)))).astype(np.float32)
        dec = off[:, None, None, None] + t * sc[:, None, None, None]
    return volnet.LatentGrid(encoding, np.ascontiguousarray(dec.astype(np.float32)), raw, off, sc)


# ------------------------------------------------------------------------------------------------ CPU stand-ins
class _Camera:  # camera.generate_rays: renderer_camera_kernels.cuh:12-43 + renderer_camera.cuh:34-52, fp32
    def __init__(self, matrix, fov_y): self.m, self.fov = matrix.astype(np.float32), np.float32(fov_y)

    def generate_rays(self, width, height, double_precision):
        eye, right, up = (torch.from_numpy(self.m[0, i]) for i in range(3))
        front = torch.linalg.cross(up, right)
        tan_y = torch.tan(torch.tensor(self.fov / 2)); tan_x = tan_y * (width / height)
        xs = 2 * (torch.arange(width, dtype=torch.float32) + 0.5) / width - 1
        ys = 2 * (torch.arange(height, dtype=torch.float32) + 0.5) / height - 1
        d = front[None, None] + xs[None, :, None] * tan_x * right[None, None] + ys[:, None, None] * tan_y * up[None, None]
        d = d / torch.linalg.norm(d, dim=-1, keepdim=True)
        return eye[None, None, None].expand(1, height, width, 3).contiguous(), d[None].contiguous()


class _PiecewiseTF:  # tf.evaluate_with_previous: renderer_tf_kernels.cuh:42-70 + renderer_tf_piecewise.cuh:27-62
    def __init__(self, tensor): self.t = torch.from_numpy(tensor[0])

    def evaluate_with_previous(self, density, dmin, dmax, previous, stepsize):
        d = density[:, 0]
        d2 = torch.clamp((d - dmin) * (1.0 / (dmax - dmin)), 0, 1)
        pos = self.t[:, 4]; R = pos.shape[0]
        # first i in [0,R-2) with pos[i+1] > d, else R-2
        gt = pos[None, 1:R - 1] > d2[:, None]
        i = torch.where(gt.any(1), gt.float().argmax(1), torch.full_like(d2, R - 2, dtype=torch.long))
        p0, p1 = pos[i], pos[i + 1]
        f = (torch.minimum(torch.maximum(d2, p0), p1) - p0) / (p1 - p0)
        c = self.t[i, :4] + f[:, None] * (self.t[i + 1, :4] - self.t[i, :4])
        c = torch.cat([c[:, :3], c[:, 3:] * stepsize], 1)
        return torch.where((d >= dmin)[:, None], c, torch.zeros_like(c))


class _Volume:
    def __init__(self, box_min, box_size): self._mn, self._sz = box_min, box_size
    def box_min(self): return sys.modules["pyrenderer"].float3(*self._mn)
    def box_size(self): return sys.modules["pyrenderer"].float3(*self._sz)


CASES = [
    # name, layers, activation, fourierstd, grid_res, outputmode, time keyframes, last-layer gain, last-layer bias shift
    # Random-init networks put all densities into a few percent of [0,1]; the output layer of the
    # reference module is rescaled in place (gain, shift) BEFORE export so that renders exercise the whole TF.
    ("c1_relu_nerf", "32:32:32", "ReLU", -1, 16, "density:direct", 0, 6.0, 0.45),
    ("c3_snakealt_nerf", "32:32:32", "SnakeAlt:2", -1, 16, "density:direct", 0, 12.0, 0.45),
    ("c3_relu_gauss", "32:32:32", "ReLU", 1.0, 16, "density:direct", 0, 6.0, 0.0),
    ("sine_48", "48:48", "Sine:2", 0.7, 8, "density", 0, 4.0, 0.0),
    ("snake_64x6", "64:64:64:64:64", "Snake:2", 0.5, 12, "density:direct", 0, 3.0, 0.4),
    ("rgbo_32", "32:32:32", "ReLU", 1.0, 8, "rgbo", 0, 1.0, 0.0),
    ("c5_time", "32:32:32", "ReLU", -1, 16, "density:direct", 3, 6.0, 0.45),
]
BOX_MIN, BOX_SIZE = (-0.5, -0.5, -0.5), (1.0, 1.0, 1.0)


def main():
    torch.set_num_threads(8)
    tf_tensor = scenes.piecewise_tf_tensor()
    np.save(os.path.join(HERE, "plume_tf.npy"), tf_tensor)
    cam = scenes.camera_reference_frame(**scenes.DEFAULT_CAMERA)
    for seed, (name, layers, activation, fstd, gres, omode, T, gain, shift) in enumerate(CASES):
        hidden = int(layers.split(':')[0])
        opt = make_opt(layers, activation, (hidden - 4) // 2, fstd, gres, outputmode=omode, time_dependent=T > 0)
        if T > 0:
            opt['time_keyframes'] = f"0:{T}:1"; opt['ensembles'] = "0:1:1"
        torch.manual_seed(1000 + seed)
        net = SceneRepresentationNetwork(opt, _FakeInput(T), torch.float32, torch.device('cpu'))
        net.eval()
        with torch.no_grad():
            last = getattr(net._hidden_layers, 'linear%d' % len(layers.split(':')))
            last.weight.mul_(gain); last.bias.add_(shift)
        vn = export(net, opt, BOX_MIN, BOX_SIZE)
        g = torch.Generator().manual_seed(7)
        pos = torch.rand((256, 3), generator=g)          # unittests/testSRN.cpp:330-340: 256 uniform positions
        time_val = 1.25 if T > 0 else 0.0
        args = [torch.zeros(256), torch.full((256,), time_val), torch.zeros(256)]  # tf, time, ensemble
        with torch.no_grad():
            world = net(pos, *args, 'world').numpy()
            screen = net(pos, *args, 'screen').numpy()
        out = dict(volnet=np.frombuffer(volnet.dumps(vn), np.uint8), positions=pos.numpy(), world=world, screen=screen,
                   time=np.float32(time_val))
        if omode.startswith("density"):   # DVR through the reference's PyTorch marcher, 48x48
            W = H = 48
            step = 1.0 / 128
            ie = types.SimpleNamespace(camera=_Camera(cam, scenes.DEFAULT_FOV_Y), volume=_Volume(BOX_MIN, BOX_SIZE),
                                       ray_evaluator=types.SimpleNamespace(tf=_PiecewiseTF(tf_tensor), min_density=0.0, max_density=1.0))
            rt = Raytracing(ie, 'density', step, W, H, torch.float32, torch.device('cpu'))
            rs, rd = ie.camera.generate_rays(W, H, False)
            nargs = [torch.zeros(1, 1), torch.full((1, 1), time_val), torch.zeros(1, 1), 'screen']  # inference.py:560-563
            with torch.no_grad():
                img = rt._full_trace_forward(net, rs, rd, ie.ray_evaluator.tf, False, False, nargs)  # (1,H,W,4)
            out.update(dvr=img.numpy()[0].transpose(2, 0, 1), dvr_size=np.int32(W), dvr_step=np.float32(step), camera=cam)
        path = os.path.join(HERE, name + ".npz")
        np.savez_compressed(path, **out)
        print(name, "->", path, os.path.getsize(path) // 1024, "KiB", "world range", world.min(), world.max())
        if name == "c3_relu_gauss":
            # byte-encoded latent grids: the same network with its grid stored as 8-bit texels. The reference
            # encodes in C++ (LatentGrid ctor, restated in encode_grid) and its PyTorch network has no byte mode. The
            # expected outputs come from the reference network with its latent fetch redirected to what the reference KERNEL
            # does with such a grid (renderer_volume_tensorcores.cuh:573-605): grid_sample on the normalised texels x/255,
            # then EncodeGridValue (offset + scale * x, resp. offset + scale * sqrt2 * erfinv((2 - 1e-4)(x - 0.5))).
            grid_param = net._volumetric_latent_space
            orig = grid_param.detach().clone()
            for enc, tag in ((volnet.GRID_BYTE_LINEAR, "bytelinear"), (volnet.GRID_BYTE_GAUSSIAN, "bytegauss")):
                eg = encode_grid(orig.numpy()[0], enc)
                vb = export(net, opt, BOX_MIN, BOX_SIZE); vb.time_grids = [eg]
                import volnet.network as refnet
                real_sample = refnet.F.grid_sample
                off_t = torch.from_numpy(eg.offset_or_mean)[None, :, None, None, None]; sc_t = torch.from_numpy(eg.scale_or_std)[None, :, None, None, None]

                def sample_then_decode(inp, grid, **kw):
                    x = real_sample(inp, grid, **kw)
                    if enc == volnet.GRID_BYTE_GAUSSIAN:
                        x = float(np.sqrt(2.0)) * torch.erfinv((2.0 - 1e-4) * (x - 0.5))
                    return off_t + sc_t * x
                with torch.no_grad():
                    grid_param.copy_(torch.from_numpy(eg.raw.astype(np.float32) / np.float32(255))[None])
                    refnet.F.grid_sample = sample_then_decode
                    try:
                        wb = net(pos, *args, 'world').numpy()
                    finally:
                        refnet.F.grid_sample = real_sample
                    grid_param.copy_(orig)
                pb = os.path.join(HERE, f"{name}_{tag}.npz")
                np.savez_compressed(pb, volnet=np.frombuffer(volnet.dumps(vb), np.uint8), positions=pos.numpy(), world=wb, screen=wb,
                                    time=np.float32(0.0), camera=cam, dvr_size=np.int32(48), dvr_step=np.float32(1.0 / 128))
                print(pb, os.path.getsize(pb) // 1024, "KiB", "max |decoded - float| grid", np.abs(eg.values - orig.numpy()[0]).max())


if __name__ == "__main__":
    main()
