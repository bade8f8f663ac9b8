"""Builder diagnostic: ms per C3 frame of each kernel mode (CUDA events, L2 flushed), incl. the reference-exact validation mode."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import fvsrn_b200  # noqa
from fvsrn_b200 import pyrenderer as pr, scenes, volnet
plume = np.load(os.path.join(ROOT, "tests", "golden", "plume_tf.npy"))
net = scenes.make_network(hidden=32, num_layers=4, grid_res=32, seed=42, activation="ReLU", last_gain=6.0)
cam = scenes.camera_reference_frame(**scenes.DEFAULT_CAMERA)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for mode in ("tcgen05", "mma_sync", "reference_exact"):
    sn = pr.SceneNetwork.from_bytes(volnet.dumps(net)); sn.kernel_mode = mode
    vol = pr.VolumeInterpolationNetwork(); vol.set_network(sn)
    ie = pr.ImageEvaluatorSimple(); ie.volume = vol
    ie.camera.set_parameters(torch.from_numpy(np.ascontiguousarray(cam, np.float32)))
    ie.ray_evaluator.tf.tensor.value = torch.from_numpy(plume).cuda(); ie.ray_evaluator.stepsize = 1 / 512
    out = ie.render(1024, 1024); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ms = []
    for _ in range(5):
        flush.zero_(); e0.record(); ie.render(1024, 1024, out); e1.record(); torch.cuda.synchronize(); ms.append(e0.elapsed_time(e1))
    print(f"C3 1024^2 step 1/512, kernel mode {mode}: {np.median(ms):.3f} ms/frame")

# normals: finite differences (mma.sync family, 7 evaluations per shaded step), the adjoint pass (reference-arithmetic family, one
# forward + one backward), and the reference's own kernels of both kinds on the same GPU
import parity  # noqa: E402
import ref_kernel  # noqa: E402
brdf = dict(flags=("magnitude", "phong"), light_type=1, light=(0.3, -0.5, -0.8), magnitude_scaling=2.0, ambient=0.3, specular=0.5,
            magnitude_center=1.0, magnitude_radius=0.8, specular_exponent=8)
for label, grad in (("finite differences, mma_sync kernels", 1), ("finite differences, reference-exact kernels", 1), ("adjoint pass, mma_sync kernels", 2),
                    ("adjoint pass, reference-exact kernels", 2)):
    sn = pr.SceneNetwork.from_bytes(volnet.dumps(net)); sn.kernel_mode = "reference_exact" if "exact" in label else "mma_sync"
    vol = pr.VolumeInterpolationNetwork(); vol.set_network(sn)
    vol.gradient_mode = pr.VolumeInterpolationNetwork.FINITE_DIFFERENCES if grad == 1 else pr.VolumeInterpolationNetwork.ADJOINT_METHOD
    ie = pr.ImageEvaluatorSimple(); ie.volume = vol
    ie.camera.set_parameters(torch.from_numpy(np.ascontiguousarray(cam, np.float32)))
    ie.ray_evaluator.tf.tensor.value = torch.from_numpy(plume).cuda(); ie.ray_evaluator.stepsize = 1 / 512
    b = ie.ray_evaluator.brdf; b.enable_magnitude_scaling = b.enable_phong = True
    b.magnitude_scaling, b.ambient, b.specular, b.magnitude_center, b.magnitude_radius, b.specular_exponent = 2.0, 0.3, 0.5, 1.0, 0.8, 8
    b.light_follows_camera = False; b.light_direction = pr.double3(0.3, -0.5, -0.8)
    out = ie.render(1024, 1024); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ms = []
    for _ in range(3):
        flush.zero_(); e0.record(); ie.render(1024, 1024, out); e1.record(); torch.cuda.synchronize(); ms.append(e0.elapsed_time(e1))
    print(f"C3 1024^2 step 1/512 with Phong shading, normals by {label}: {np.median(ms):.3f} ms/frame")
if ref_kernel.available():
    for variant in ("h32x4_relu_fd_phong", "h32x4_relu_adjoint_phong"):
        _, _, ms = parity.render_reference(net, 1024, 1024, cam, scenes.DEFAULT_FOV_Y, plume, 1 / 512, variant=variant, brdf=dict(brdf, light_type=0 if "fd" in variant else 1),
                                           fd_stepsize=1 / 256, timing_iters=3)
        print(f"C3 1024^2 step 1/512 with Phong shading, the reference's kernel {variant}: {ms:.3f} ms/frame")
