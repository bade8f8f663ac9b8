"""Builder diagnostic (not a test): error statistics of our kernels vs the reference kernel for selected configs."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)

import fvsrn_b200  # noqa
from fvsrn_b200 import scenes, volnet
import parity
import oracle
import torch_ref

plume = np.load(os.path.join(ROOT, "tests", "golden", "plume_tf.npy"))
from test_gpu_baseline_configs import CASES, make_case_network

names = sys.argv[1].split(",")
size_override = int(sys.argv[2]) if len(sys.argv) > 2 else 0
for name in names:
    kw, size, step, tf_kind, time = CASES[name][:5]
    absorption = CASES[name][5] if len(CASES[name]) > 5 else 1.0
    if size_override:
        size = size_override
    net = make_case_network(kw)
    cam = scenes.camera_reference_frame(**scenes.DEFAULT_CAMERA)
    tfp = plume.copy(); tfp[..., 3] *= absorption
    tf = scenes.texture_tf_from_piecewise(tfp, 256) if tf_kind == 1 else tfp
    want, wsteps, _ = parity.render_reference(net, size, size, cam, scenes.DEFAULT_FOV_Y, tf, step, tf_kind=tf_kind, time=0.0 if time is None else time)
    print(f"== {name} size {size}: reference alpha max {want[:,3].max():.4f} mean {want[:,3].mean():.4f} steps mean {wsteps.mean():.1f} finite {np.isfinite(want[:,:4]).all()}")
    data = volnet.dumps(net)
    for mode in ("tcgen05", "mma_sync"):
        img, steps = parity.render_ours(data, mode, size, size, cam, tf, step, tf_kind=tf_kind, time=time)
        r = parity.compare(img, steps, want, wsteps)
        d = np.abs(img[0, :4] - want[0, :4]).max(0)
        y, x = np.unravel_index(np.argmax(d), d.shape)
        print(f"  {mode}: alpha max {img[:,3].max():.4f} mean {img[:,3].mean():.4f} | {r}")
        print(f"    worst pixel ({x},{y}): ours {img[0,:4,y,x]} ref {want[0,:4,y,x]} steps {steps[0,y,x]} vs {wsteps[0,y,x]}; pixels >2e-3: {(d>2e-3).sum()} >5e-3: {(d>5e-3).sum()}")
        # CPU oracles on a small crop around the worst pixel
        x0, y0 = max(0, min(size - 8, x - 4)), max(0, min(size - 8, y - 4))
        for label, kwo in (("oracle fp16acc+tex (reference arithmetic)", dict(tex_filter=True, acc32=False)),
                           ("oracle fp32acc+tex (tcgen05 arithmetic)", dict(tex_filter=True, acc32=True))):
            o, _, _ = oracle.render_dvr(net, width=size, height=size, camera=cam, fov_y=scenes.DEFAULT_FOV_Y, stepsize=step, tf=tf,
                                        tf_kind=tf_kind, time=0.0 if time is None else time, tile=(x0, y0, 8, 8), **kwo)
            oc = o[0, :4, y0:y0 + 8, x0:x0 + 8]
            print(f"    {label}: vs ref {np.abs(oc - want[0,:4,y0:y0+8,x0:x0+8]).max():.2e}  vs ours {np.abs(oc - img[0,:4,y0:y0+8,x0:x0+8]).max():.2e}")
        tn = torch_ref.TorchNetwork(net)
        t, _, _ = torch_ref.render_dvr(tn, width=size, height=size, camera=cam, fov_y=scenes.DEFAULT_FOV_Y, stepsize=step, tf=tf, tile=(x0, y0, 8, 8)) if tf_kind == 0 and time is None else (None, 0, 0)
        if t is not None:
            tc = np.asarray(t)[:4]
            print(f"    torch fp32 path: vs ref {np.abs(tc - want[0,:4,y0:y0+8,x0:x0+8]).max():.2e}  vs ours {np.abs(tc - img[0,:4,y0:y0+8,x0:x0+8]).max():.2e}")
