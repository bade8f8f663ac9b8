"""Builder diagnostic: is the reference-exact kernel mode bit-identical to the reference kernel? (prints, no asserts)"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import fvsrn_b200  # noqa
from fvsrn_b200 import scenes, volnet
import parity
from test_gpu_baseline_configs import CASES, make_case_network

plume = np.load(os.path.join(ROOT, "tests", "golden", "plume_tf.npy"))
names = sys.argv[1].split(",") if len(sys.argv) > 1 else list(CASES)
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
    for mode in (sys.argv[3].split(",") if len(sys.argv) > 3 else ["reference_exact"]):
        img, steps = parity.render_ours(volnet.dumps(net), mode, size, size, cam, tf, step, tf_kind=tf_kind, time=time)
        a, b = np.nan_to_num(img, nan=-7.0), np.nan_to_num(want, nan=-7.0)
        neq = (a[0, :4] != b[0, :4]).any(0)
        print(f"{name:28s} {size:5d} {mode:16s} rgba bit-equal: {not neq.any()} ({int(neq.sum())} px differ, max-abs {np.abs(a[:, :4]-b[:, :4]).max():.3e}); "
              f"depth equal: {np.array_equal(a[0,7], b[0,7])}; steps equal: {np.array_equal(steps, wsteps)} ({int((steps != wsteps).sum())} rays differ); "
              f"saturated {float((want[:,3] >= 1-1e-5).mean()):.3f}")
        if neq.any():
            ys, xs = np.nonzero(neq)
            for y, x in list(zip(ys, xs))[:3]:
                print(f"     px ({x},{y}): ours {img[0,:4,y,x]} ref {want[0,:4,y,x]} steps {steps[0,y,x]} vs {wsteps[0,y,x]}")
