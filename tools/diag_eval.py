"""Builder diagnostic (not a test): the world-space batch query (C2) of the current library - time per 2^24 random positions and
max-abs against the CPU oracle (texture-filter weights, fp32 accumulation = the arithmetic of the tcgen05 family).
  python tools/diag_eval.py [time] [lib.so ...]   - with libraries given, each is copied over the product library in turn (A/B on one box)"""
import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "fv-srn_b200", "libfvsrn_b200.so")
libs = [a for a in sys.argv[1:] if a.endswith(".so")]
if libs:        # one fresh process per library
    keep = LIB + ".keep"
    shutil.copy(LIB, keep)
    try:
        for rep in range(2):
            for l in libs:
                shutil.copy(os.path.join(ROOT, "ab", l), LIB)
                out = subprocess.run([sys.executable, __file__] + [a for a in sys.argv[1:] if not a.endswith(".so")], capture_output=True, text=True)
                print(l, (out.stdout.strip() or out.stderr.strip()[-400:]))
    finally:
        shutil.move(keep, LIB)
    sys.exit(0)

import numpy as np
import torch
for p in (ROOT, os.path.join(ROOT, "oracle")):
    sys.path.insert(0, p)
import fvsrn_b200  # noqa
from fvsrn_b200 import pyrenderer as pr, scenes, volnet
import oracle

kw = dict(hidden=32, num_layers=4, grid_res=32, seed=42, activation="ReLU", last_gain=6.0)
timed = "time" in sys.argv[1:]
if timed:
    kw["time_keyframes"] = 4
net = scenes.make_network(**kw)
sn = pr.SceneNetwork.from_bytes(volnet.dumps(net)); sn.kernel_mode = os.environ.get("FVSRN_KERNEL_MODE", "tcgen05")
if timed:
    sn.set_time_and_ensemble(1.37, 0)
vol = pr.VolumeInterpolationNetwork(); vol.set_network(sn)
g = torch.Generator(device="cuda"); g.manual_seed(1)
pos = torch.rand((1 << 24, 3), device="cuda", generator=g)
out = vol.evaluate(pos)
torch.cuda.synchronize()
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
ms = []
for _ in range(10):
    flush.zero_()
    e0.record(); out = vol.evaluate(pos); e1.record(); torch.cuda.synchronize()
    ms.append(e0.elapsed_time(e1))
n = 1 << 15
want = oracle.evaluate(net, pos[:n].cpu().numpy(), time=1.37 if timed else 0.0, tex_filter=True, acc32=True)
d = np.abs(out[:n].cpu().numpy() - want)
print(f"C2 {sn.kernel_mode}{' two keyframes' if timed else ''}: {np.median(ms):.4f} ms per 2^24 positions = {(1 << 24) / np.median(ms) * 1e3:.3e} samples/s; "
      f"max-abs vs oracle {d.max():.2e}, mean {d.mean():.2e}")
