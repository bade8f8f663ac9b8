"""
torch_ref.py — CPU PyTorch restatement of the reference's PyTorch path (TEST / BASELINE INFRASTRUCTURE ONLY).

The reference's only CPU-runnable implementation of the hot path is its PyTorch code:
  * SceneRepresentationNetwork.forward        applications/volnet/network.py:998-1096
      InputParametrization (Fourier: cos/sin of B·x, network.py:88-167), grid_sample latent grid
      (align_corners=False, padding_mode='border', :1079-1084), nn.Linear stack + activations, OutputParametrization
      (:204-235)
  * Raytracing._full_trace_forward            applications/volnet/raytracing.py:275-329
      (+ intersection_aabb :79-92, _predict :119-155, _blend :157-166)
  * the three pyrenderer calls it makes: camera.generate_rays (renderer/renderer_camera.cuh:34-52),
    tf.evaluate_with_previous (renderer/renderer_tf_kernels.cuh:42-70 + renderer_tf_piecewise.cuh:27-62), box.
It cannot travel to the GPU box (/root/reference does not exist there), so this file restates it with plain torch ops
on a parsed .volnet. Pinned against the committed fixtures in tests/golden/ (tests/test_torch_ref.py).
Used by bench.py's ``cpu_baseline`` and ``--impl reference`` legs ("CPU torch ref" of BASELINE.json's metric).
May only be imported by tests/, __graft_entry__.smoke() and bench.py.
"""
from __future__ import annotations

import math

import numpy as np
import torch
import torch.nn.functional as F


class TorchNetwork:
    """fp32 torch model of a parsed .volnet (volnet.SceneNetwork), evaluated like network.py:998-1096."""

    def __init__(self, net, time: float = 0.0):
        assert net.num_fourier > 0
        self.net = net
        self.H = net.input_channels_out()
        self.nf = net.num_fourier
        self.has_dir = bool(net.has_direction)
        self.has_time = bool(net.has_time)
        self.B = torch.from_numpy(np.asarray(net.fourier, np.float16).astype(np.float32))  # (nf, cin), already x 2pi
        self.layers = [(torch.from_numpy(np.asarray(l.weights, np.float16).astype(np.float32)),
                        torch.from_numpy(np.asarray(l.bias, np.float16).astype(np.float32))) for l in net.layers]
        self.activation = net.layers[0].activation
        self.param = float(net.layers[0].activation_param)
        self.output_mode = net.output_mode
        self.grids = None
        self.time_value = 0.0
        if net.has_latent_grid and net.time_grids:
            # the product path stores the grid in fp16; every checker sees the same fp16-rounded grid
            g = np.stack([tg.values.astype(np.float16).astype(np.float32) for tg in net.time_grids])
            self.grids = torch.from_numpy(g)                              # (T,C,Z,Y,X)
        self.set_time(time)

    def set_time(self, time: float):
        n = self.net
        if self.grids is None:
            return
        t = min(max((time - n.time_min) / n.time_step, 0.0), float(n.time_num - 1))
        self.time_value = t
        lo = min(int(math.floor(t)), n.time_num - 1)
        hi = min(lo + 1, n.time_num - 1)
        self._lo, self._hi, self._f = lo, hi, t - lo

    def _act(self, v):
        a, p = self.activation, self.param
        if a == "ReLU": return torch.relu(v)
        if a == "Sine": return torch.sin(p * v)
        if a == "Snake": return v + (1.0 / p) * torch.sin(p * v) ** 2
        if a == "SnakeAlt": return (v + 1 - torch.cos(2 * p * v)) / (2 * p)
        if a == "None": return v
        raise ValueError(a)

    def forward(self, x: torch.Tensor, direction: torch.Tensor | None = None, mode: str = "world") -> torch.Tensor:
        """x (N,3) in the unit cube -> (N,C)."""
        N = x.shape[0]
        cols = [x, torch.full((N, 1), self.time_value if self.has_time else 0.0)]
        fin = x
        if self.has_dir:
            cols += [direction, torch.zeros((N, 1))]
            if self.net.direction_in_fourier:
                fin = torch.cat([x, direction], 1)
        ph = fin @ self.B.t()
        cols += [torch.cos(ph), torch.sin(ph)]
        if self.grids is not None:
            gp = (x * 2 - 1).view(1, N, 1, 1, 3)
            lat = F.grid_sample(self.grids[self._lo:self._lo + 1], gp, align_corners=False, padding_mode="border")[0, :, :, 0, 0].t()
            if self._hi != self._lo and self._f > 0:
                hi = F.grid_sample(self.grids[self._hi:self._hi + 1], gp, align_corners=False, padding_mode="border")[0, :, :, 0, 0].t()
                lat = (1 - self._f) * lat + self._f * hi
            cols.append(lat)
        y = torch.cat(cols, 1)
        for i, (w, b) in enumerate(self.layers):
            y = F.linear(y, w, b)
            if i + 1 < len(self.layers):
                y = self._act(y)
        m = self.output_mode
        if m == "density": return torch.sigmoid(y)
        if m == "density:direct": return torch.clamp(y, 0, 1) if mode == "screen" else y
        if m == "rgbo": return torch.cat([torch.sigmoid(y[:, :3]), F.softplus(y[:, 3:])], 1)
        if m == "rgbo:direct":
            return torch.cat([torch.clamp(y[:, :3], 0, 1), torch.clamp(y[:, 3:], min=0)], 1) if mode == "screen" else y
        raise ValueError(m)


def generate_rays(camera: np.ndarray, fov_y: float, width: int, height: int, x0=0, y0=0, w=None, h=None):
    """camera.generate_rays for a sub-rectangle: (eye (3,), dirs (h*w,3))."""
    w = width - x0 if w is None else w
    h = height - y0 if h is None else h
    m = torch.from_numpy(np.asarray(camera, np.float32).reshape(3, 3))
    eye, right, up = m[0], m[1], m[2]
    front = torch.linalg.cross(up, right)
    tan_y = math.tan(fov_y / 2); tan_x = tan_y * (width / height)
    xs = 2 * (torch.arange(x0, x0 + w, dtype=torch.float32) + 0.5) / width - 1
    ys = 2 * (torch.arange(y0, y0 + h, dtype=torch.float32) + 0.5) / height - 1
    d = front[None, None] + xs[None, :, None] * tan_x * right[None, None] + ys[:, None, None] * tan_y * up[None, None]
    d = d / torch.linalg.norm(d, dim=-1, keepdim=True)
    return eye, d.reshape(-1, 3)


def tf_piecewise(tf: torch.Tensor, d: torch.Tensor, dmin: float, dmax: float, stepsize: float) -> torch.Tensor:
    """tf (R,5), d (N,) -> (N,4) rgb + absorption*stepsize; zero where d < dmin (renderer_tf_kernels.cuh:28-31)."""
    d2 = torch.clamp((d - dmin) * (1.0 / (dmax - dmin)), 0, 1)
    pos = tf[:, 4]; R = pos.shape[0]
    gt = pos[None, 1:R - 1] > d2[:, None]
    if R > 2:
        i = torch.where(gt.any(1), gt.float().argmax(1), torch.full_like(d2, R - 2, dtype=torch.long))
    else:           # two control points: one interval
        i = torch.zeros_like(d2, dtype=torch.long)
    p0, p1 = pos[i], pos[i + 1]
    f = (torch.minimum(torch.maximum(d2, p0), p1) - p0) / (p1 - p0)
    c = tf[i, :4] + f[:, None] * (tf[i + 1, :4] - tf[i, :4])
    c = torch.cat([c[:, :3], c[:, 3:] * stepsize], 1)
    return torch.where((d >= dmin)[:, None], c, torch.zeros_like(c))


@torch.no_grad()
def render_dvr(tnet: TorchNetwork, *, width, height, camera, fov_y, stepsize, tf, density_min=0.0, density_max=1.0,
               tile=None, max_steps_override=None, alpha_early_out=1.0 - 1e-5):
    """Raytracing._full_trace_forward (raytracing.py:275-329) for one view (optionally one tile of it).
    Returns (rgba (4,h,w) float32, executed network samples, useful samples). "useful" counts the samples a marcher with
    early termination needs: t < tmax and alpha < alpha_early_out before the step (the unit bench.py's metric uses)."""
    n = tnet.net
    x0, y0, w, h = tile if tile is not None else (0, 0, width, height)
    eye, dirs = generate_rays(camera, fov_y, width, height, x0, y0, w, h)
    box_min = torch.tensor(n.box_min, dtype=torch.float32); box_size = torch.tensor(n.box_size, dtype=torch.float32)
    inv = 1.0 / dirs
    t135 = (box_min - eye) * inv; t246 = (box_min + box_size - eye) * inv
    tmin = torch.max(torch.minimum(t135, t246), dim=1, keepdim=True)[0]
    tmax = torch.min(torch.maximum(t135, t246), dim=1, keepdim=True)[0]
    max_steps = int(torch.max(tmax - tmin).item() / stepsize) if max_steps_override is None else int(max_steps_override)
    tft = torch.from_numpy(np.asarray(tf, np.float32).reshape(-1, 5)) if tf is not None else None
    color = torch.zeros((dirs.shape[0], 3)); alpha = torch.zeros((dirs.shape[0], 1))
    is_rgbo = n.output_mode.startswith("rgbo")
    useful = 0
    for t in range(max_steps):
        tcur = tmin + t * stepsize
        useful += int(((tcur < tmax) & (alpha < alpha_early_out)).sum())
        vol = (eye + tcur * dirs - box_min) / box_size
        pred = tnet.forward(vol, dirs if tnet.has_dir else None, "screen")
        if is_rgbo:
            c = torch.cat([pred[:, :3], pred[:, 3:] * stepsize], 1)
        else:
            c = tf_piecewise(tft, pred[:, 0], density_min, density_max, stepsize)
        a = torch.where(tcur < tmax, 1 - torch.exp(-c[:, 3:]), torch.zeros(1, 1))
        color = color + (1 - alpha) * c[:, :3] * a
        alpha = alpha + (1 - alpha) * a
    rgba = torch.cat([color, alpha], 1).view(h, w, 4).permute(2, 0, 1).contiguous().numpy()
    return rgba, max_steps * dirs.shape[0], useful
