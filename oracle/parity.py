"""
TEST INFRASTRUCTURE — parity of the B200 kernels against the REFERENCE's own tensor-core kernel at the BASELINE configs.

The reference kernel (oracle/_ref/*.ptx, built by oracle/build_ref_ptx.py from the unmodified renderer/*.cuh) and the
product kernels render the same frame on the same GPU from the same .volnet weights, fp16-rounded latent grid, camera
and transfer function; this module returns the numbers north_star gates: RGBA max-abs (<= 2e-3), PSNR (>= 50 dB) and
the fraction of rays whose early-termination step count differs (reference kernel patched, test-only, to export it).
Used by tests/test_gpu_baseline_configs.py and by bench.py's `parity` sub-object. Nothing here is product code.
"""
import numpy as np

RGBA_TOL = 2e-3
PSNR_MIN = 50.0


def psnr(a, b):
    mse = float(np.mean((np.asarray(a, np.float64) - np.asarray(b, np.float64)) ** 2))
    return 200.0 if mse == 0 else 10.0 * np.log10(1.0 / mse)


def render_ours(net_bytes, mode, width, height, camera, tf, stepsize, tf_kind=0, time=None, tile=None, spp=1,
                fd_stepsize=None, brdf=None, blend=None, early_out=None, density_range=None, adjoint=False):
    """One frame through the product path (pyrenderer mirror -> C ABI) -> (img (B,8,H,W) numpy, steps (B,H,W) numpy).
    fd_stepsize: finite-difference normals with that step; brdf: dict as ref_kernel.ReferenceDvrKernel.render takes it plus
    "flags" (subset of "magnitude", "phong") and "light_type" (0 point / 1 directional)."""
    import torch
    import fvsrn_b200  # noqa: F401
    from fvsrn_b200 import pyrenderer as pr
    sn = pr.SceneNetwork.from_bytes(net_bytes); sn.kernel_mode = mode
    if time is not None:
        sn.set_time_and_ensemble(float(time), 0)
    vol = pr.VolumeInterpolationNetwork(); vol.set_network(sn)
    ie = pr.ImageEvaluatorSimple(); ie.volume = vol
    ie.camera.set_parameters(torch.from_numpy(np.ascontiguousarray(camera, np.float32)))
    if tf_kind == 2:        # identity TF: tf = (scale_absorption, scale_emission)
        ie.ray_evaluator.tf = pr.TransferFunctionIdentity(float(tf[0]), float(tf[1]))
    else:
        if tf_kind == 1:
            ie.ray_evaluator.tf = pr.TransferFunctionTexture()
        ie.ray_evaluator.tf.tensor.value = torch.from_numpy(np.ascontiguousarray(tf, np.float32)).cuda()
    ie.ray_evaluator.stepsize = stepsize
    ie.samples_per_pixel = spp
    if fd_stepsize is not None:
        vol.gradient_mode = pr.VolumeInterpolationNetwork.FINITE_DIFFERENCES
        vol.finite_differences_stepsize = fd_stepsize
    if adjoint:             # GRADIENT_MODE 2: backward pass through the network, latent-grid step 1 / (4 * resolution)
        vol.gradient_mode = pr.VolumeInterpolationNetwork.ADJOINT_METHOD
    if brdf is not None:
        b = ie.ray_evaluator.brdf
        b.enable_magnitude_scaling, b.enable_phong = "magnitude" in brdf["flags"], "phong" in brdf["flags"]
        b.magnitude_scaling, b.ambient, b.specular = brdf["magnitude_scaling"], brdf["ambient"], brdf["specular"]
        b.magnitude_center, b.magnitude_radius, b.specular_exponent = brdf["magnitude_center"], brdf["magnitude_radius"], brdf["specular_exponent"]
        b.light_follows_camera = False
        b.light_type = pr.BRDFLambert.Point if brdf["light_type"] == 0 else pr.BRDFLambert.Directional
        b.light_position = b.light_direction = pr.double3(*brdf["light"])
    if blend is not None:
        ie.ray_evaluator.blending.blendMode = pr.Blending.Alpha if blend == "alpha" else pr.Blending.BeerLambert
    if early_out is not None:
        ie.ray_evaluator.early_out = bool(early_out)
    if density_range is not None:
        ie.ray_evaluator.min_density, ie.ray_evaluator.max_density = density_range
    ie.collect_statistics = True
    ie.tile = tile
    img = ie.render(width, height).cpu().numpy()
    return img, ie.last_step_counts.cpu().numpy()


def render_reference(net, width, height, camera, fov_y, tf, stepsize, tf_kind=0, time=0.0, timing_iters=0, variant=None,
                     fd_stepsize=0.01, brdf=None, density_range=(0.0, 1.0), steps_variant=None):
    """The same frame with the reference's kernel -> (rgba8 image, step counts, ms or None). variant: a named build of
    oracle/build_ref_ptx.py (the gradient / shading kernels; they have no step-count build: steps come back as None)."""
    import ref_kernel
    tfname = {1: "texture", 2: "identity"}.get(tf_kind, "piecewise")
    rk = ref_kernel.ReferenceDvrKernel(net, time=time, tf=tfname, variant=variant, fd_stepsize=fd_stepsize)
    ms = None
    if timing_iters > 0:
        want, ms = rk.render(width, height, camera, fov_y, stepsize, tf, timing_iters=timing_iters, brdf=brdf)
    else:
        want = rk.render(width, height, camera, fov_y, stepsize, tf, brdf=brdf, density_min=density_range[0], density_max=density_range[1])
    want = want.cpu().numpy()
    if variant is not None and steps_variant is None:
        return want, None, ms
    if steps_variant is not None:
        counted = ref_kernel.ReferenceDvrKernel(net, variant=steps_variant, time=time, tf=tfname).render(
            width, height, camera, fov_y, stepsize, tf, density_min=density_range[0], density_max=density_range[1]).cpu().numpy()
        assert np.array_equal(want[:, :4], counted[:, :4]), "the step-count patch must not touch the image"
        return want, np.rint(counted[:, 4]).astype(np.int32), ms
    counted = ref_kernel.ReferenceDvrKernel(net, patched_steps=True, time=time, tf=tfname).render(
        width, height, camera, fov_y, stepsize, tf).cpu().numpy()
    assert np.array_equal(want[:, :4], counted[:, :4]), "the step-count patch must not touch the image"
    return want, np.rint(counted[:, 4]).astype(np.int32), ms


def compare(img, steps, want, want_steps):
    """-> dict with the gated numbers."""
    d = np.abs(img[:, :4] - want[:, :4])
    mism = steps != want_steps
    return {"rgba_max_abs": float(d.max()), "psnr_db": float(psnr(img[:, :4], want[:, :4])),
            "step_count_mismatch_frac": float(mism.mean()), "step_count_max_diff": int(np.abs(steps - want_steps).max()),
            "rays": int(mism.size), "total_steps_ref": int(want_steps.sum()), "total_steps": int(steps.sum())}
