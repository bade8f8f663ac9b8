"""
TEST INFRASTRUCTURE — builds the REFERENCE's own fused DVR kernel (and nothing of ours) into PTX.

The reference assembles its hot path at run time with NVRTC (renderer/kernel_loader.cpp:197-273, 368-463): a list of
#defines + #includes of renderer/*.cuh, compiled for compute_XX, loaded with cuModuleLoadDataEx. This script repeats that
recipe where /root/reference exists (the build container), reading the headers from where they lie and
writing ONLY into oracle/_ref/: one PTX per network configuration plus a JSON with the lowered symbol names. The PTX
travels to the GPU box (oracle/_ref is git-ignored, not gpurun-ignored), where oracle/ref_kernel.py loads it through the
driver API, packs the reference's __constant__ blocks and launches ImageEvaluatorSimpleKernel — the
reference's tensor-core kernel on the B200, used by tests/ as a parity comparand and by bench.py as a GPU baseline.
No reference source is copied into the repository.

The host library of the reference cannot be built here; the two cudAD headers it would pull in are absent
submodules, so two shim headers stand in for them (forward declarations only — the DVR forward path uses none of it).
"""
import glob
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_ref")
REF = os.environ.get("FVSRN_REFERENCE", "/root/reference")

SHIM_FORWARD = r"""
#pragma once
#define CUDAD_CALL_NOCONSTEXPR __host__ __device__ inline
namespace std { using ::exp; }
namespace cudAD {
template<typename T, int N> struct fvar;
template<int N, typename V, typename I> __host__ __device__ fvar<float2, N> make_float2in(const V&, const I&);
template<int N, typename V, typename I> __host__ __device__ fvar<float3, N> make_float3in(const V&, const I&);
template<int N, typename V, typename I> __host__ __device__ fvar<float4, N> make_float4in(const V&, const I&);
template<int N, typename V, typename I> __host__ __device__ fvar<double2, N> make_double2in(const V&, const I&);
template<int N, typename V, typename I> __host__ __device__ fvar<double3, N> make_double3in(const V&, const I&);
template<int N, typename V, typename I> __host__ __device__ fvar<double4, N> make_double4in(const V&, const I&);
}
"""

NAMES = ["ImageEvaluatorSimpleKernel", "&cameraReferenceFrameParameters", "&volumeInterpolationTensorcoresParameters",
         "&rayEvaluationSteppingDvrParameters", "&brdfLambertParameters"]
TF_NAMES = {"piecewise": "&transferFunctionPiecewiseParameters", "texture": "&transferFunctionTextureParameters",
            "identity": "&transferFunctionIdentityParameters"}
TF_TYPES = {"piecewise": ("::kernel::TransferFunctionPiecewise", "renderer_tf_piecewise.cuh"),
            "texture": ("::kernel::TransferFunctionTexture", "renderer_tf_texture.cuh"),
            "identity": ("::kernel::TransferFunctionIdentity", "renderer_tf_identity.cuh")}


def grid_source(early_out=1, blend_mode=1, tf="piecewise", interpolation=1):
    """The same DVR kernel over the ground-truth density grid: VolumeInterpolationGrid::getDefines for a .cvol density feature
    (volume_interpolation_grid.cpp:913-993: texture path, trilinear, new grid-resolution behaviour), per-thread tracing."""
    d = [
        ("KERNEL_DOUBLE_PRECISION", 0), ("KERNEL_SYNCHRONIZED_TRACING", 0),
        ("VOLUME_INTERPOLATION_GRID__INTERPOLATION", interpolation), ("VOLUME_INTERPOLATION_GRID__TEXTURE_TYPE", "float"),
        ("VOLUME_INTERPOLATION_GRID__TEXTURE_EXTRACTOR", "getDirect"), ("VOLUME_INTERPOLATION_GRID__GRID_RESOLUTION_OLD_BEHAVIOR", 0),
        ("RAY_EVALUATION_STEPPING__VOLUME_INTERPOLATION_T", "::kernel::VolumeInterpolationGrid"),
        ("RAY_EVALUATION_STEPPING__TRANSFER_FUNCTION_T", TF_TYPES[tf][0]),
        ("RAY_EVALUATION_STEPPING__BLENDING_T", "::kernel::Blending"),
        ("RAY_EVALUATION_STEPPING__BRDF_T", "::kernel::BRDFLambert"),
        ("RAY_EVALUATION_STEPPING__ENABLE_EARLY_OUT", early_out), ("RAY_EVALUATION_STEPPING__SKIP_TRANSFER_FUNCTION", 0),
        ("BLENDING_MODE", blend_mode), ("BRDF_LAMBERT_LIGHT_TYPE", 1),
    ]
    src = "".join(f"#define {k} {v}\n" for k, v in d)
    for h in ("renderer_camera.cuh", "renderer_volume_grid.cuh", TF_TYPES[tf][1], "renderer_blending.cuh",
              "renderer_brdf_lambert.cuh", "renderer_ray_evaluation_stepping_dvr.cuh"):
        src += f'#include "{h}"\n'
    src += "#define IMAGE_EVALUATOR__CAMERA_T ::kernel::CameraReferenceFrame\n"
    src += "#define IMAGE_EVALUATOR__RAY_EVALUATOR_T ::kernel::RayEvaluationSteppingDvr\n"
    src += '#include "renderer_image_evaluator_simple.cuh"\n'
    return src


def main_source(hidden16, num_hidden, nf, activation, output_mode, block_size, early_out=1, blend_mode=1, first_last_shared=0,
                encoding=0, tf="piecewise", gradient_mode=0, brdf=(), light_type=1, steps_variant=True, skip_tf=0,
                use_direction=0, pass_time=0, grid16=1):
    """The translation unit IImageEvaluator::getKernel + ImageEvaluatorSimple::render assemble.
    encoding: LATENT_GRID_ENCODING 0 float / 1 byte-linear / 2 byte-gaussian (volume_interpolation_network.cpp:1211);
    tf "texture": TransferFunctionTexture in its tensor form (transfer_function_texture.cpp: TRANSFER_FUNCTION_TEXTURE__USE_TENSOR,
    pre-integration off). gradient_mode: GRADIENT_MODE 0 off-or-predicted / 1 finite differences;
    brdf: subset of ("magnitude", "phong") = BRDFLambert::getDefines (brdf.cpp:372-394); light_type 0 point / 1 directional."""
    d = [
        ("KERNEL_DOUBLE_PRECISION", 0), ("KERNEL_SYNCHRONIZED_TRACING", 1), ("BLOCK_SIZE", block_size),
        ("NUM_HIDDEN_LAYERS", num_hidden), ("HIDDEN_CHANNELS_DIV16", hidden16), ("HAS_FOURIER_FEATURES", 1),
        ("NUM_FOURIER_FEATURES", nf), ("USE_DIRECTION", use_direction), ("ACTIVATION", activation), ("OUTPUT_MODE", output_mode),
        ("FIRST_AND_LAST_IN_SHARED_MEMORY", first_last_shared), ("LATENT_GRID_CHANNELS_DIV16", grid16), ("LATENT_GRID_ENCODING", encoding),
        ("PASS_TIME_TO_NETWORK", pass_time), ("GRADIENT_MODE", gradient_mode),
        ("RAY_EVALUATION_STEPPING__VOLUME_INTERPOLATION_T", "::kernel::VolumeInterpolationTensorcores"),
        ("RAY_EVALUATION_STEPPING__TRANSFER_FUNCTION_T", TF_TYPES[tf][0]),
        ("RAY_EVALUATION_STEPPING__BLENDING_T", "::kernel::Blending"),
        ("RAY_EVALUATION_STEPPING__BRDF_T", "::kernel::BRDFLambert"),
        ("RAY_EVALUATION_STEPPING__ENABLE_EARLY_OUT", early_out), ("RAY_EVALUATION_STEPPING__SKIP_TRANSFER_FUNCTION", skip_tf),
        ("BLENDING_MODE", blend_mode), ("BRDF_LAMBERT_LIGHT_TYPE", light_type),
    ]
    if "magnitude" in brdf:
        d.append(("BRDF_LAMBERT_ENABLE_MAGNITUDE_SCALING", ""))
    if "phong" in brdf:
        d.append(("BRDF_LAMBERT_ENABLE_PHONG", ""))
    if tf == "texture":
        d += [("TRANSFER_FUNCTION_TEXTURE__USE_TENSOR", 1), ("TRANSFER_FUNCTION_TEXTURE__PREINTEGRATION_MODE", 0)]
    src = "".join(f"#define {k} {v}\n" for k, v in d)
    for h in ("renderer_camera.cuh", "renderer_volume_tensorcores.cuh", TF_TYPES[tf][1], "renderer_blending.cuh",
              "renderer_brdf_lambert.cuh", "renderer_ray_evaluation_stepping_dvr.cuh"):
        src += f'#include "{h}"\n'
    src += "#define IMAGE_EVALUATOR__CAMERA_T ::kernel::CameraReferenceFrame\n"
    src += "#define IMAGE_EVALUATOR__RAY_EVALUATOR_T ::kernel::RayEvaluationSteppingDvr\n"
    src += '#include "renderer_image_evaluator_simple.cuh"\n'
    return src


def eval_source(hidden16, num_hidden, nf, activation, output_mode, block_size, encoding=0, channels=1, **_):
    """The translation unit IVolumeInterpolation::evaluate assembles for the world-space batch query (volume_interpolation.cpp:
    26-127): the network defines of VolumeInterpolationNetwork::getDefines + EvaluateNoBatches (renderer_volume_kernels1.cuh)."""
    d = [
        ("KERNEL_DOUBLE_PRECISION", 0), ("KERNEL_SYNCHRONIZED_TRACING", 1), ("BLOCK_SIZE", block_size),
        ("NUM_HIDDEN_LAYERS", num_hidden), ("HIDDEN_CHANNELS_DIV16", hidden16), ("HAS_FOURIER_FEATURES", 1),
        ("NUM_FOURIER_FEATURES", nf), ("USE_DIRECTION", 0), ("ACTIVATION", activation), ("OUTPUT_MODE", output_mode),
        ("FIRST_AND_LAST_IN_SHARED_MEMORY", 0), ("LATENT_GRID_CHANNELS_DIV16", 1), ("LATENT_GRID_ENCODING", encoding),
        ("PASS_TIME_TO_NETWORK", 0), ("GRADIENT_MODE", 0),
    ]
    src = "".join(f"#define {k} {v}\n" for k, v in d)
    src += '#include "renderer_volume_tensorcores.cuh"\n'
    src += "#define VOLUME_INTERPOLATION_T ::kernel::VolumeInterpolationTensorcores\n#define OUTPUT_CHANNELS %d\n#define VOLUME_USE_DIRECTION 0\n" % channels
    src += '#include "renderer_volume_kernels1.cuh"\n'
    return src


# TEST-ONLY patch, applied to the in-memory copy of two headers for the "*_steps" variants (never written anywhere): the
# reference kernel does not output its per-ray step count (terminationIndex is a local, ..stepping_dvr.cuh:83-84,156), so
# the count of valid steps is routed through the (otherwise zero) normal.x channel. Arithmetic and control flow of the
# march are untouched. The unpatched PTX is the one used for RGBA parity and timing.
STEP_PATCHES = {
    "renderer_ray_evaluation_stepping_dvr.cuh": [
        ("int terminationIndex;", "int terminationIndex; int fvsrnValidSteps = 0;"),
        ("#if KERNEL_SYNCHRONIZED_TRACING==0\n\t\t\t\tif (!isValid) break;", "fvsrnValidSteps += isValid;\n#if KERNEL_SYNCHRONIZED_TRACING==0\n\t\t\t\tif (!isValid) break;"),
        ("\t\t\treturn output;\n\t\t}\n\t};", "\t\t\toutput.normal = make_real3(real_t(fvsrnValidSteps), 0, 0);\n\t\t\treturn output;\n\t\t}\n\t};"),
    ],
    "renderer_image_evaluator_simple.cuh": [
        ("out.normal += nout.normal * nout.color.w;", "out.normal += nout.normal;"),
    ],
}


def compile_ptx(src, patch_steps=False, tf="piecewise", volume_symbol="&volumeInterpolationTensorcoresParameters", names_override=None):
    from cuda.bindings import nvrtc

    def ck(r):
        if r[0] != nvrtc.nvrtcResult.NVRTC_SUCCESS:
            raise RuntimeError(str(r[0]))
        return r[1:] if len(r) > 2 else (r[1] if len(r) == 2 else None)

    headers, names = [], []
    for f in sorted(glob.glob(os.path.join(REF, "renderer", "*.cuh"))):
        text = open(f, "rb").read()
        if patch_steps and os.path.basename(f) in STEP_PATCHES:
            t = text.decode().replace("\r\n", "\n")
            for old, new in STEP_PATCHES[os.path.basename(f)]:
                assert t.count(old) == 1, (os.path.basename(f), old, t.count(old))
                t = t.replace(old, new)
            text = t.encode()
        headers.append(text); names.append(os.path.basename(f).encode())
    # the two cudAD headers renderer_cudad_bridge.cuh includes by name (:7-8); the submodule is absent -> shims
    headers += [SHIM_FORWARD.encode(), b"#pragma once\n"]
    names += [b"forward.h", b"forward_vector.h"]
    prog = ck(nvrtc.nvrtcCreateProgram(src.encode(), b"main.cu", len(headers), headers, names))
    names_all = names_override or ([n if n != "&volumeInterpolationTensorcoresParameters" else volume_symbol for n in NAMES] + [TF_NAMES[tf]])
    for n in names_all:
        ck(nvrtc.nvrtcAddNameExpression(prog, n.encode()))
    cuda_inc = os.path.join(os.environ.get("CUDA_HOME", "/usr/local/cuda"), "include")
    opts = [b"--std=c++17", b"--use_fast_math", b"--generate-line-info", b"--gpu-architecture=compute_75",
            b"-I" + cuda_inc.encode(), b"-D__NVCC__=1", b"-DCUDA_NO_HOST=1", b"-DUSE_DOUBLE_PRECISION=0"]
    res = nvrtc.nvrtcCompileProgram(prog, len(opts), opts)
    if res[0] != nvrtc.nvrtcResult.NVRTC_SUCCESS:
        n = ck(nvrtc.nvrtcGetProgramLogSize(prog)); log = b" " * n
        ck(nvrtc.nvrtcGetProgramLog(prog, log))
        raise RuntimeError("NVRTC failed:\n" + log.decode(errors="replace")[-4000:])
    n = ck(nvrtc.nvrtcGetPTXSize(prog)); ptx = b" " * n
    ck(nvrtc.nvrtcGetPTX(prog, ptx))
    lowered = {nm: ck(nvrtc.nvrtcGetLoweredName(prog, nm.encode())).decode() for nm in names_all}
    return ptx.rstrip(b"\x00 "), lowered


CONFIGS = {
    # name: hidden16, num_hidden (linear layers - 2), nf, activation, output_mode (1 = density:direct), block size
    "h32x4_relu": dict(hidden16=2, num_hidden=2, nf=14, activation="ReLU", output_mode=1, block_size=256),
    "h32x4_snakealt": dict(hidden16=2, num_hidden=2, nf=14, activation="SnakeAlt", output_mode=1, block_size=256),
    # other widths of the reference's own network sweep (eval_NetworkConfigsGrid.py:35-37); block size from computeMaxWarps
    "h48x4_relu": dict(hidden16=3, num_hidden=2, nf=22, activation="ReLU", output_mode=1, block_size=256),
    "h96x3_relu": dict(hidden16=6, num_hidden=1, nf=46, activation="ReLU", output_mode=1, block_size=32),
    "h128x2_relu": dict(hidden16=8, num_hidden=0, nf=62, activation="ReLU", output_mode=1, block_size=32),
    # alpha blending (BLENDING_MODE 0) without early termination: the other values of the two marcher defines
    "h32x4_relu_alpha_noearly": dict(hidden16=2, num_hidden=2, nf=14, activation="ReLU", output_mode=1, block_size=256, blend_mode=0, early_out=0),
    # network inputs beyond the position: ray direction (USE_DIRECTION 1), direction also inside the Fourier features (2), the
    # keyframe time (PASS_TIME_TO_NETWORK); 32 latent channels (two 16-channel groups)
    "h32x4_relu_dir": dict(hidden16=2, num_hidden=2, nf=12, activation="ReLU", output_mode=1, block_size=256, use_direction=1),
    "h32x4_relu_dirfourier": dict(hidden16=2, num_hidden=2, nf=12, activation="ReLU", output_mode=1, block_size=256, use_direction=2),
    "h32x4_relu_time": dict(hidden16=2, num_hidden=2, nf=14, activation="ReLU", output_mode=1, block_size=256, pass_time=1),
    "h32x4_relu_grid32": dict(hidden16=2, num_hidden=2, nf=14, activation="ReLU", output_mode=1, block_size=256, grid16=2),
    "h32x4_sine": dict(hidden16=2, num_hidden=2, nf=14, activation="Sine", output_mode=1, block_size=256),
    "h32x4_snake": dict(hidden16=2, num_hidden=2, nf=14, activation="Snake", output_mode=1, block_size=256),
    # BASELINE C4 network: 64x6, 30 Fourier features; computeMaxWarps (volume_interpolation_network.cpp:987-1041) leaves
    # (48 KiB - 43648 B weights) / (64*2*32 B per warp) = 1 warp per block
    "h64x6_relu": dict(hidden16=4, num_hidden=4, nf=30, activation="ReLU", output_mode=1, block_size=32),
    # texture transfer function (tensor form), byte-encoded latent grids
    "h32x4_relu_identitytf": dict(hidden16=2, num_hidden=2, nf=14, activation="ReLU", output_mode=1, block_size=256, tf="identity"),
    "h32x4_relu_textf": dict(hidden16=2, num_hidden=2, nf=14, activation="ReLU", output_mode=1, block_size=256, tf="texture"),
    "h32x4_relu_bytelinear": dict(hidden16=2, num_hidden=2, nf=14, activation="ReLU", output_mode=1, block_size=256, encoding=1),
    "h32x4_relu_bytegauss": dict(hidden16=2, num_hidden=2, nf=14, activation="ReLU", output_mode=1, block_size=256, encoding=2),
    # networks that predict the gradient next to the density (4 = densitygrad, 6 = densitygrad:cubic), finite
    # differences on a density network, Phong shading and gradient-magnitude scaling of BRDFLambert. The step-count patch
    # uses the normal channel, so these are built unpatched only.
    "h32x4_relu_densgrad": dict(hidden16=2, num_hidden=2, nf=14, activation="ReLU", output_mode=4, block_size=256, steps_variant=False),
    "h32x4_relu_densgrad_cubic_phong": dict(hidden16=2, num_hidden=2, nf=14, activation="ReLU", output_mode=6, block_size=256,
                                            brdf=("magnitude", "phong"), light_type=1, steps_variant=False),
    # (colour networks - OUTPUT_MODE rgbo, SKIP_TRANSFER_FUNCTION 1 - do NOT compile in the reference with the tensor-core volume:
    # renderer_ray_evaluation_stepping_dvr.cuh:109 calls volume.evalNormal(position, rayDir, batch), an overload
    # VolumeInterpolationTensorcores does not have. Their batch query does: see the *_rgbo_eval build below.)
    "h32x4_relu_fd_phong": dict(hidden16=2, num_hidden=2, nf=14, activation="ReLU", output_mode=1, block_size=256, gradient_mode=1,
                                brdf=("magnitude", "phong"), light_type=0, steps_variant=False),
    # GRADIENT_MODE 2: the adjoint pass through the network (renderer_volume_tensorcores.cuh:1205-1530); BLOCK_SIZE 128 because
    # its activation storage needs (NUM_HIDDEN_LAYERS + 1) * BLOCK_SIZE * HIDDEN_CHANNELS halves more shared memory
    "h32x4_relu_adjoint_phong": dict(hidden16=2, num_hidden=2, nf=14, activation="ReLU", output_mode=1, block_size=128, gradient_mode=2,
                                     brdf=("magnitude", "phong"), light_type=1, steps_variant=False),
    "h32x4_relu_densgrad_adjoint": dict(hidden16=2, num_hidden=2, nf=14, activation="ReLU", output_mode=4, block_size=128, gradient_mode=2,
                                        brdf=("magnitude", "phong"), light_type=1, steps_variant=False),
    "h32x4_sine_adjoint": dict(hidden16=2, num_hidden=2, nf=14, activation="Sine", output_mode=1, block_size=128, gradient_mode=2, steps_variant=False),
    "h32x4_snake_adjoint": dict(hidden16=2, num_hidden=2, nf=14, activation="Snake", output_mode=1, block_size=128, gradient_mode=2, steps_variant=False),
    "h32x4_relu_grid32_adjoint": dict(hidden16=2, num_hidden=2, nf=14, activation="ReLU", output_mode=1, block_size=128, gradient_mode=2, grid16=2,
                                      steps_variant=False),
    "h48x4_relu_adjoint": dict(hidden16=3, num_hidden=2, nf=22, activation="ReLU", output_mode=1, block_size=64, gradient_mode=2, steps_variant=False),
    "h32x4_snakealt_adjoint_phong": dict(hidden16=2, num_hidden=2, nf=14, activation="SnakeAlt", output_mode=0, block_size=128, gradient_mode=2,
                                         brdf=("magnitude", "phong"), light_type=0, steps_variant=False),
}
# (time-interpolated grids, C5, need no build of their own: cLatentGridB != A and cLatentGridInterpolation = t are
# constants of the same h32x4 kernels, renderer_volume_tensorcores.cuh:590-604)


def main():
    if not os.path.isdir(os.path.join(REF, "renderer")):
        print(f"build_ref_ptx: {REF} not present - keeping the prebuilt files in {OUT}")
        return 0
    os.makedirs(OUT, exist_ok=True)
    meta = {}
    only = set(sys.argv[1:])
    for base, cfg in CONFIGS.items():
        if only and base not in only:
            continue
        for name, patch in ((base, False), (base + "_steps", True)):
            if patch and not cfg.get("steps_variant", True):
                continue
            ptx, lowered = compile_ptx(main_source(**cfg), patch_steps=patch, tf=cfg.get("tf", "piecewise"))
            open(os.path.join(OUT, name + ".ptx"), "wb").write(ptx)
            meta[name] = dict(cfg, lowered=lowered, ptx=name + ".ptx", step_count_patch=patch)
            print(f"build_ref_ptx: {name}: {len(ptx)} bytes of PTX, kernel {lowered['ImageEvaluatorSimpleKernel']}")
    # the world-space batch query (BASELINE C2) of the two 32x4 networks: EvaluateNoBatches
    for base, name, extra in (("h32x4_relu", "h32x4_relu_eval", {}), ("h32x4_snakealt", "h32x4_snakealt_eval", {}),
                              ("h32x4_relu", "h32x4_sine_eval", dict(activation="Sine")), ("h32x4_relu", "h32x4_snake_eval", dict(activation="Snake")),
                              ("h32x4_relu", "h32x4_relu_rgbo_eval", dict(output_mode=2, channels=4)),
                              ("h32x4_relu", "h32x4_relu_rgbo_direct_eval", dict(output_mode=3, channels=4))):
        if only and name not in only:
            continue
        cfg = dict(CONFIGS[base], **extra)
        ptx, lowered = compile_ptx(eval_source(**cfg), names_override=["EvaluateNoBatches", "&volumeInterpolationTensorcoresParameters"])
        open(os.path.join(OUT, name + ".ptx"), "wb").write(ptx)
        meta[name] = dict(cfg, eval=True, lowered=lowered, ptx=name + ".ptx", step_count_patch=False)
        print(f"build_ref_ptx: {name}: {len(ptx)} bytes of PTX, kernel {lowered['EvaluateNoBatches']}")
    for name, patch in (("grid_trilinear", False), ("grid_trilinear_steps", True)):
        if len(sys.argv) > 1 and "grid_trilinear" not in sys.argv[1:]:
            continue
        ptx, lowered = compile_ptx(grid_source(), patch_steps=patch, volume_symbol="&volumeInterpolationGridParameters")
        open(os.path.join(OUT, name + ".ptx"), "wb").write(ptx)
        meta[name] = dict(grid=True, lowered=lowered, ptx=name + ".ptx", step_count_patch=patch)
        print(f"build_ref_ptx: {name}: {len(ptx)} bytes of PTX")
    path = os.path.join(OUT, "ref_kernels.json")
    if len(sys.argv) > 1 and os.path.exists(path):      # partial rebuild: keep the other entries
        meta = dict(json.load(open(path)), **meta)
    json.dump(meta, open(path, "w"), indent=1)
    return 0


if __name__ == "__main__":
    sys.exit(main())
