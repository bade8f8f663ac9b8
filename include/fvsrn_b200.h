/*
 * fvsrn_b200.h — C ABI of the B200-native fV-SRN hot path (libfvsrn_b200.so).
 *
 * One hot path, behind the reference's own module API:
 *   the fused scene-representation-network (SRN) evaluation inside the DVR ray-marcher.
 *
 * Every entry point replaces (or is what a maintainer would bind from) one reference
 * interface; the file:line cited beside each is the reference code it stands in for
 * (paths relative to the fV-SRN tree).  No torch / pybind types cross this boundary:
 * plain pointers, sizes, an opaque handle and a CUDA stream passed as void*.
 *
 * Conventions
 *   - every function returns 0 on success, non-zero on failure; the message is in
 *     fvsrn_last_error() (thread-local).  The reference throws std::runtime_error
 *     (renderer/pytorch_utils.h:5-13, kernel_loader.cpp:19-31); a C++/pybind shim maps
 *     non-zero -> throw, the Python host mirror maps it to RuntimeError.
 *   - device pointers are suffixed _dev, host pointers _host.
 *   - `stream` is a cudaStream_t (the caller's current stream, as
 *     renderer/iimage_evaluator.cpp:167-170 does); NULL = default stream.
 *   - there is NO CPU fallback: compute entry points fail when no CUDA device is present.
 */
#ifndef FVSRN_B200_H
#define FVSRN_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FVSRN_ABI_VERSION 3

typedef struct fvsrn_network fvsrn_network; /* opaque: renderer::SceneNetwork + its GPU resources */

/* Layer::Activation, renderer/volume_interpolation_network.h (names: volume_interpolation_network.cpp:223-230) */
enum fvsrn_activation {
    FVSRN_ACT_RELU = 0, FVSRN_ACT_SINE = 1, FVSRN_ACT_SNAKE = 2, FVSRN_ACT_SNAKEALT = 3,
    FVSRN_ACT_SIGMOID = 4, FVSRN_ACT_NONE = 5
};
/* OutputParametrization::OutputMode, volume_interpolation_network.cpp:165-175 */
enum fvsrn_output_mode {
    FVSRN_OUT_DENSITY = 0, FVSRN_OUT_DENSITY_DIRECT = 1, FVSRN_OUT_RGBO = 2, FVSRN_OUT_RGBO_DIRECT = 3,
    /* the network predicts the gradient next to the density (4 last-layer outputs; ":cubic" stores its cube
     * root) or gradient + curvature (6 outputs; the DVR path uses density and gradient), renderer_volume_tensorcores.cuh:1073-1127 */
    FVSRN_OUT_DENSITY_GRADIENT = 4, FVSRN_OUT_DENSITY_GRADIENT_DIRECT = 5, FVSRN_OUT_DENSITY_GRADIENT_CUBIC = 6,
    FVSRN_OUT_DENSITY_CURVATURE = 7, FVSRN_OUT_DENSITY_CURVATURE_DIRECT = 8
};
/* SceneNetwork::GradientMode, volume_interpolation_network.h / renderer_volume_tensorcores.cuh:132-134, 1166-1203 */
enum fvsrn_gradient_mode {
    FVSRN_GRADIENT_OFF_OR_DIRECT = 0,       /* no normals, or the ones the network predicts (densitygrad* / densitycurvature*) */
    FVSRN_GRADIENT_FINITE_DIFFERENCES = 1,  /* central differences: 6 more SRN evaluations per shaded step */
    /* the backward pass through the SRN (:1205-1530): d density / d position from one more pass with transposed weights,
     * the latent grid differentiated by central differences of its texture (step 1 / (4 * grid resolution) by default).
     * Fast kernel modes: the mma.sync family's own adjoint pass; FVSRN_KERNEL_REFERENCE_EXACT: bit-identical to the reference
     * kernel built with GRADIENT_MODE 2. Density / densitygrad* networks without direction input, float latent grids. */
    FVSRN_GRADIENT_ADJOINT_METHOD = 2
};
/* BRDFLambert::LightType as the kernel sees it, renderer_brdf_lambert.cuh:6-7 */
enum fvsrn_light_type { FVSRN_LIGHT_POINT = 0, FVSRN_LIGHT_DIRECTIONAL = 1 };
enum { FVSRN_BRDF_MAGNITUDE_SCALING = 1, FVSRN_BRDF_PHONG = 2 };
/* LatentGrid::Encoding, volume_interpolation_network.cpp:564-614 */
enum fvsrn_grid_encoding { FVSRN_GRID_FLOAT = 0, FVSRN_GRID_BYTE_LINEAR = 1, FVSRN_GRID_BYTE_GAUSSIAN = 2 };
/* BLENDING_MODE, renderer/renderer_blending.cuh:7-8 */
enum fvsrn_blend_mode { FVSRN_BLEND_ALPHA = 0, FVSRN_BLEND_BEER_LAMBERT = 1 };
/* transfer function kinds on the hot path */
enum fvsrn_tf_kind {
    FVSRN_TF_PIECEWISE = 0, /* renderer/renderer_tf_piecewise.cuh:27-62, tensor (B,R,5) = r,g,b,absorption,pos */
    FVSRN_TF_TEXTURE = 1,   /* renderer/renderer_tf_texture.cuh:41-60 (tensor variant), (B,R,4) rgba */
    FVSRN_TF_IDENTITY = 2   /* renderer/renderer_tf_identity.cuh: rgb = density*scale_c, absorption = density*scale_a */
};

/* IImageEvaluator::ChannelMode, renderer/iimage_evaluator.h:19-26 */
enum fvsrn_channel_mode { FVSRN_CHANNEL_MASK = 0, FVSRN_CHANNEL_NORMAL = 1, FVSRN_CHANNEL_DEPTH = 2, FVSRN_CHANNEL_COLOR = 3 };
/* what a DVR launch (or fvsrn_extract_color_ex) writes:
 *   RAW8_F32  the reference's (B,8,H,W) fp32 tensor: rgb, alpha, normal xyz, depth (renderer_image_evaluator_simple.cuh:106-124)
 *   RGBA_F32  (B,4,H,W) fp32: the selected channel, what extract_color returns (iimage_evaluator.cpp:26-135)
 *   RGBA8     (B,H,W) uint32 = rgbaToInt (renderer_utils.cuh:48-57): the reference's display format (SelectOutputChannelKernel /
 *             TonemappingKernel writing the GL texture, iimage_evaluator_cuda.cu:21-38, 168-186)
 *   RGBA_F16  (B,H,W) 4 x fp16 (B200 addition: half the bytes of RGBA_F32 for the multi-GPU gather) */
enum fvsrn_output_format { FVSRN_OUTPUT_RAW8_F32 = 0, FVSRN_OUTPUT_RGBA_F32 = 1, FVSRN_OUTPUT_RGBA8 = 2, FVSRN_OUTPUT_RGBA_F16 = 3 };

/* ------------------------------------------------------------------------------------------
 * Network lifetime and .volnet I/O
 *   replaces renderer::SceneNetwork::{load,save}   volume_interpolation_network.cpp:1059-1104
 * ------------------------------------------------------------------------------------------ */
int  fvsrn_abi_version(void);
const char* fvsrn_last_error(void);

int  fvsrn_network_create(fvsrn_network** out);                                   /* SceneNetwork()  :797-804 */
void fvsrn_network_free(fvsrn_network* net);
int  fvsrn_network_load_volnet(const void* bytes, size_t n, fvsrn_network** out); /* SceneNetwork::load :1059-1086 */
/* two-call pattern: buf==NULL -> *written = required size */
int  fvsrn_network_save_volnet(const fvsrn_network* net, void* buf, size_t cap, size_t* written); /* ::save :1088-1104 */

/* Builders mirroring the pyrenderer surface (volume_interpolation_network.cpp:1828-1972) */
int  fvsrn_network_set_input(fvsrn_network* net, int has_time, int has_direction);
/* B is (nf, cin) row-major fp32, cin = 3 or 6; premultiplied=0 multiplies by 2*pi.
 * InputParametrization::setFourierMatrixFromTensor :129-156 */
int  fvsrn_network_set_fourier(fvsrn_network* net, const float* B_host, int nf, int cin, int premultiplied);
int  fvsrn_network_disable_fourier(fvsrn_network* net);
int  fvsrn_network_set_output_mode(fvsrn_network* net, const char* mode);          /* OutputModeFromString :177-184 */
/* W is (cout, cin) row-major fp32 (torch nn.Linear.weight), b is (cout).
 * SceneNetwork::addLayerFromTorch :896-921 (+ addLayer re-layout :806-894) */
int  fvsrn_network_add_layer(fvsrn_network* net, const float* W_host, const float* b_host,
                             int cout, int cin, const char* activation, float activation_param);
/* LatentGridTimeAndEnsemble(time_min,time_num,time_step,ensemble_min,ensemble_num) :756-796 */
int  fvsrn_network_set_latent_grid_layout(fvsrn_network* net, int time_min, int time_num, int time_step,
                                          int ensemble_min, int ensemble_num);
/* grid is (C,Z,Y,X) fp32 row-major (torch (1,C,Z,Y,X)); LatentGrid(t, encoding) :290-334.
 * returns the encoding error (mean abs) through *encoding_error if non-NULL. */
int  fvsrn_network_set_time_grid(fvsrn_network* net, int index, const float* grid_host,
                                 int C, int Z, int Y, int X, int encoding, double* encoding_error);
int  fvsrn_network_set_ensemble_grid(fvsrn_network* net, int index, const float* grid_host,
                                     int C, int Z, int Y, int X, int encoding, double* encoding_error);
int  fvsrn_network_set_box(fvsrn_network* net, const float box_min[3], const float box_size[3]);
int  fvsrn_network_get_box(const fvsrn_network* net, float box_min[3], float box_size[3]);
int  fvsrn_network_valid(const fvsrn_network* net);                                /* SceneNetwork::valid :940-985; 1 = valid */
int  fvsrn_network_num_parameters(const fvsrn_network* net);                       /* :1043-1055 */
int  fvsrn_network_compute_max_warps(const fvsrn_network* net, int only_shared, int adjoint); /* :987-1041 */
int  fvsrn_network_set_time_and_ensemble(fvsrn_network* net, float time, int ensemble);       /* :923-938 */
/* VolumeInterpolationNetwork::{gradient_mode, finite_differences_stepsize} (volume_interpolation_network.cpp:1976-1995):
 * how the DVR frame obtains normals; stepsize in world units (reference default 1/256). Density-type networks only.
 * With FVSRN_GRADIENT_ADJOINT_METHOD the third argument is the central-difference step through the latent grid in normalised
 * grid coordinates, i.e. 1 / (resolution * adjoint_latent_grid_central_differences_stepsize_scale) of the reference (:1809-1812);
 * <= 0 selects the reference default 1 / (resolution * 4). */
int  fvsrn_network_set_gradient_mode(fvsrn_network* net, int gradient_mode, float finite_differences_stepsize);

/* Introspection (what a binding needs to expose .input/.output/.latent_grid read-only) */
typedef struct fvsrn_network_info {
    int has_time, has_direction, num_fourier, direction_in_fourier;
    int output_mode;          /* fvsrn_output_mode (or 4..8) */
    int num_layers;           /* linear layers incl. the last */
    int hidden_channels;      /* H */
    int latent_channels;      /* total latent-grid channels (0 if none) */
    int grid_x, grid_y, grid_z;
    int time_min, time_num, time_step, ensemble_min, ensemble_num;
    int grid_encoding;
    int activation;           /* of the hidden layers */
    float activation_param;
    int channels_out;         /* 1 (density*) or 4 (rgbo*) */
    int gradient_mode;        /* fvsrn_gradient_mode */
    float finite_differences_stepsize;
} fvsrn_network_info;
int  fvsrn_network_get_info(const fvsrn_network* net, fvsrn_network_info* info);

/* Which sm_100a kernel family evaluates the SRN (no reference analogue; the reference picks wmma by NVRTC defines):
 *   FVSRN_KERNEL_TCGEN05  (default) tcgen05.mma with TMEM accumulators (fp16 for ReLU networks, fp32 otherwise), 128 rays per CTA in lock-step
 *   FVSRN_KERNEL_MMA_SYNC reference-precision mode: mma.sync m16n8k16 with fp16 accumulation in the reference's
 *                         K order (renderer_volume_tensorcores.cuh:856-910), 32 rays per warp
 *   FVSRN_KERNEL_REFERENCE_EXACT  validation mode (DVR frame only): every fp16 and fp32 operation of the reference's
 *                         compiled kernel in the reference's order - tex3D<float4> latent fetch, hsin/hcos/__hdiv of
 *                         cuda_fp16.hpp, fp16-accumulate HMMA with the bias preloaded, sequential fp16 last layer, fast-math
 *                         fp32 ray setup / TF / blending - so that images and early-termination step counts are
 *                         bit-identical to ImageEvaluatorSimpleKernel on the same GPU (csrc/kernels_exact.cuh). Density-type
 *                         networks with a latent grid, with or without the time input; the ray direction as input and 32
 *                         latent channels for hidden widths 32 and 64. */
enum { FVSRN_KERNEL_MMA_SYNC = 0, FVSRN_KERNEL_TCGEN05 = 1, FVSRN_KERNEL_REFERENCE_EXACT = 2 };
int  fvsrn_network_set_kernel_mode(fvsrn_network* net, int mode);
int  fvsrn_network_get_kernel_mode(const fvsrn_network* net);

/* Upload weights (fragment-ordered fp16) and the fp16 channel-last latent grids to `device`.
 * replaces SceneNetwork::fillConstantMemory :1236-1417, LatentGrid::copyGridToGPU :524-535 and the
 * per-network NVRTC specialisation in renderer/kernel_loader.cpp:197-273 (kernels are AOT sm_100a). */
int  fvsrn_network_prepare(fvsrn_network* net, int device, void* stream);
int  fvsrn_network_clear_gpu_resources(fvsrn_network* net);                        /* :1106-1113 */

/* ------------------------------------------------------------------------------------------
 * World-space batch query
 *   replaces IVolumeInterpolation::evaluate + EvaluateNoBatches
 *   renderer/volume_interpolation.cpp:26-127, renderer/renderer_volume_kernels1.cuh:15-71
 *   positions (N,3) fp32 in [0,1]^3 (the reference forces the box to the unit cube, :45-48),
 *   directions (N,3) or NULL, out (N,channels_out) fp32.
 * ------------------------------------------------------------------------------------------ */
int  fvsrn_evaluate(fvsrn_network* net, const float* positions_dev, const float* directions_dev,
                    float* out_dev, int64_t n, void* stream);
/* same call with HOST buffers (copies inside; pinned memory makes them asynchronous) */
int  fvsrn_evaluate_host(fvsrn_network* net, const float* positions_host, const float* directions_host,
                         float* out_host, int64_t n, void* stream);

/* ------------------------------------------------------------------------------------------
 * DVR frame
 *   replaces ImageEvaluatorSimple::render -> ImageEvaluatorSimpleKernel
 *   renderer/image_evaluator_simple.cpp:198-361, renderer/renderer_image_evaluator_simple.cuh:36-127,
 *   renderer/renderer_ray_evaluation_stepping_dvr.cuh:49-157, renderer/renderer_camera.cuh:31-52,
 *   renderer/renderer_tf_piecewise.cuh:27-62, renderer/renderer_tf_texture.cuh:41-60,
 *   renderer/renderer_blending.cuh:36-51
 * ------------------------------------------------------------------------------------------ */
typedef struct fvsrn_dvr_params {
    int width, height, batch;       /* full image size and camera batch B */
    /* sub-rectangle actually rendered (screen-tile sharding); ndc is still relative to width x height.
     * tile_w/tile_h == 0 means the full image. */
    int tile_x0, tile_y0, tile_w, tile_h;
    /* camera: (B,3,3) fp32 rows = eye, right, up (camera.cpp:495-549) */
    const float* camera_matrix_dev;
    float fov_y_radians;
    /* RayEvaluationSteppingDvrParameters, renderer_ray_evaluation_stepping_dvr.cuh:22-29 */
    float stepsize;
    float alpha_early_out;          /* 1 - 1e-5 in the reference (ray_evaluation_stepping.cpp:603-623) */
    float density_min, density_max;
    int   early_out;                /* RAY_EVALUATION_STEPPING__ENABLE_EARLY_OUT */
    int   blend_mode;               /* fvsrn_blend_mode */
    /* transfer function */
    int   tf_kind;                  /* fvsrn_tf_kind */
    const float* tf_dev;            /* piecewise: (tf_batch,R,5); texture: (tf_batch,R,4); identity: 2 floats */
    int   tf_resolution;            /* R */
    int   tf_batch;                 /* 1 (broadcast) or B */
    /* samples per pixel: 1 = reference DVR (ray_evaluation.h:61); n>1 = deterministic stratified
     * sqrt(n) x sqrt(n) sub-pixel offsets (no reference analogue) */
    int   samples_per_pixel;
    /* per-view latent time (view-sharded time sweeps): NULL = use set_time_and_ensemble for all views,
     * else host array of `batch` absolute times */
    const float* view_times_host;
    /* output strides in ELEMENTS of the (B,8,H,W) fp32 tensor (any layout, renderer/pytorch_utils.h:39-43) */
    int64_t out_stride_b, out_stride_c, out_stride_y, out_stride_x;
    /* optional debug/statistics outputs (may be NULL) */
    int32_t* step_count_dev;        /* (B,H,W) int32 contiguous: valid steps of each ray (last sample of the pixel) */
    unsigned long long* counters_dev; /* [0] += valid samples, [1] += executed (lock-step) samples */
    /* Fused colour extraction: what every caller runs right after render (inference.py:621) happens in the
     * pixel write instead of a second pass over a 32 B/pixel tensor. All zero = the reference's raw tensor.
     *   output_format != RAW8_F32: `out` is the selected channel in that format; the out_stride_* are in elements of it
     *   (fp32 values for RGBA_F32 with out_stride_c between the 4 channels; uint32 / half4 pixels for RGBA8 / RGBA_F16,
     *   out_stride_c unused). channel_mode: COLOR (with optional tonemapping at a FIXED exposure, as with the reference's
     *   fix_max_exposure), MASK or NORMAL; DEPTH needs the frame's min/max depth first -> fvsrn_extract_color_ex. */
    int   output_format;            /* fvsrn_output_format */
    int   channel_mode;             /* fvsrn_channel_mode; 0-initialised params with output_format 0 ignore it */
    int   tonemap;                  /* ImageEvaluatorSimple::useTonemapping_ */
    float max_exposure;             /* ImageEvaluatorSimple::getExposureForTonemapping(), image_evaluator_simple.cpp:363-369 */
    /* Shading: BRDFLambert with scalar parameters, renderer_brdf_lambert.cuh:19-103, brdf.cpp:124-135, 372-452.
     * brdf_flags = 0 is the reference default (colour passes through). Both options need normals: a network that predicts
     * them or fvsrn_network_set_gradient_mode. light = position (point) or direction (directional); a binding fills it from
     * the camera when light_follows_camera is set (brdf.cpp:478-497). */
    int   brdf_flags;               /* FVSRN_BRDF_MAGNITUDE_SCALING | FVSRN_BRDF_PHONG */
    int   light_type;               /* fvsrn_light_type */
    float light[3];
    float brdf_magnitude_scaling, brdf_ambient, brdf_specular, brdf_magnitude_center, brdf_magnitude_radius;
    int   brdf_specular_exponent;
} fvsrn_dvr_params;

int  fvsrn_render_dvr(fvsrn_network* net, const fvsrn_dvr_params* p, float* out_dev, void* stream);
/* HOST-buffer variant: camera matrix, TF and output image live on the host; the call stages them
 * (H2D camera+TF, D2H image) on `stream` and synchronises before returning. `p->camera_matrix_dev`
 * and `p->tf_dev` are interpreted as HOST pointers, `out_host` is (B,8,H,W) with the given strides. */
int  fvsrn_render_dvr_host(fvsrn_network* net, const fvsrn_dvr_params* p, float* out_host, void* stream);

/* ------------------------------------------------------------------------------------------
 * (B,8,H,W) -> (B,4,H,W) colour extraction; replaces IImageEvaluator::extractColor /
 * SelectOutputChannelKernel, renderer/iimage_evaluator_cuda.cu:21-80 (ChannelColor only). SURVEY f1.
 * ------------------------------------------------------------------------------------------ */
int  fvsrn_extract_color(const float* img_dev, float* out_dev, int batch, int height, int width, void* stream);
/* The full IImageEvaluator::ExtractColor (iimage_evaluator.cpp:26-135) on a contiguous (B,8,H,W) fp32 tensor: channel
 * selection (mask / normal / depth / colour), optional tonemapping of the colour channel (TonemappingKernel2,
 * iimage_evaluator_cuda.cu:217-260), output as (B,4,H,W) fp32, packed RGBA8 or half4 (fvsrn_output_format, not RAW8).
 * ChannelDepth normalises by the frame's min / max depth like the reference (a torch min/max there: NaN depth of empty
 * pixels propagates). */
int  fvsrn_extract_color_ex(const float* img_dev, void* out_dev, int batch, int height, int width,
                            int channel_mode, int tonemap, float max_exposure, int output_format, void* stream);

/* fvsrn_extract_color_ex without any allocation. No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same arguments and results as fvsrn_extract_color_ex, which takes the scratch of its ChannelDepth min / max reduction from
 * cudaMallocAsync; here that scratch is a static __device__ array inside the library (192 KiB, held by every GPU that loads the
 * library until exit), and the call makes no cudaMalloc / cudaFree / allocating library call.
 *   - limit: batch * height * width <= FVSRN_NOALLOC_MAX_PIXELS (2^26, e.g. 16 frames of 2048^2); larger inputs are
 *     rejected with an error, not truncated;
 *   - ALL calls share the one array: calls that overlap on the GPU (different streams or host threads) race on it - keep
 *     them on one stream or order them with events. Nothing carries over between calls: each call writes every word it reads. */
#define FVSRN_NOALLOC_MAX_PIXELS (1 << 26)
int  fvsrn_extract_color_ex_noalloc(const float* img_dev, void* out_dev, int batch, int height, int width,
                                    int channel_mode, int tonemap, float max_exposure, int output_format, void* stream);

/* ------------------------------------------------------------------------------------------
 * Tensor operators beside the fused path: what the reference's PyTorch marcher (applications/volnet/
 * raytracing.py:103, 150-154) calls per frame / per step, so that it can run against this library on the GPU as well.
 * ------------------------------------------------------------------------------------------ */
/* ICamera::generateRays -> CameraGenerateRayKernel, renderer/renderer_camera_kernels.cuh:12-45, camera.cpp:204-209:
 * camera (B,3,3) -> ray origins and unit directions, both (B,H,W,3) fp32 contiguous */
int  fvsrn_generate_rays(const float* camera_matrix_dev, int batch, float fov_y_radians, int width, int height,
                         float* ray_start_dev, float* ray_dir_dev, void* stream);
/* ITransferFunction::evaluate / evaluate_with_previous -> EvaluateTF*, renderer/renderer_tf_kernels.cuh:11-70,
 * transfer_function.cpp:123-152: densities (N) -> rgb + absorption*stepsize (N,4); batch 0 of the TF tensor; stepsize = 1
 * for plain evaluate (the previous density only enters pre-integrated texture TFs, which are off the path) */
int  fvsrn_evaluate_tf(int tf_kind, const float* tf_dev, int tf_resolution, const float* densities_dev, float* colors_dev,
                       int64_t n, float density_min, float density_max, float stepsize, void* stream);

/* ------------------------------------------------------------------------------------------
 * Ground-truth volume: the .cvol container and the density grid rendered through the SAME marcher - what the
 * reference's eval scripts compare the SRN renders against.
 *   replaces renderer::Volume (load / save, renderer/volume.cpp:278-384, 618-800; format renderer/volume.h:42-77),
 *   VolumeInterpolationGrid (texture path, trilinear; renderer/renderer_volume_grid.cuh:110-139, 196-231; box
 *   = +-worldSize/2, volume_interpolation_grid.cpp:193-198) and ImageEvaluatorSimple::render with that volume.
 * ------------------------------------------------------------------------------------------ */
typedef struct fvsrn_volume fvsrn_volume;   /* opaque: renderer::Volume + the 3-D texture of one feature */
typedef struct fvsrn_volume_info {
    int num_features;
    int feature;                /* the feature the fields below describe */
    int size_x, size_y, size_z, channels;
    int data_type;              /* Volume::DataType: 0 uchar, 1 ushort, 2 float */
    float world_size[3];
    float box_min[3], box_size[3];
} fvsrn_volume_info;
int  fvsrn_volume_load_cvol(const void* bytes, size_t n, fvsrn_volume** out);      /* "CVOL" v1 and legacy "cvol", raw or LZ4 */
int  fvsrn_volume_save_cvol(const fvsrn_volume* vol, void* buf, size_t cap, size_t* written);  /* two-call pattern; uncompressed */
/* density (Z,Y,X) fp32, x fastest (Volume::MipmapLevel::idx) */
int  fvsrn_volume_from_dense(const float* density_host, int size_x, int size_y, int size_z, const float world_size[3], fvsrn_volume** out);
int  fvsrn_volume_get_info(const fvsrn_volume* vol, int feature, fvsrn_volume_info* info);
void fvsrn_volume_free(fvsrn_volume* vol);
/* upload one (1-channel) feature as a 3-D texture: MipmapLevel::copyCpuToGpu, volume.cpp:120-166 */
int  fvsrn_volume_prepare(fvsrn_volume* vol, int device, int feature, void* stream);
/* the DVR frame of the prepared feature; same parameters, output formats and step / sample counters as fvsrn_render_dvr */
int  fvsrn_render_dvr_grid(fvsrn_volume* vol, const fvsrn_dvr_params* p, float* out_dev, void* stream);
/* IVolumeInterpolation::evaluate for the grid: positions (N,3) in [0,1]^3 -> densities (N) */
int  fvsrn_evaluate_grid(fvsrn_volume* vol, const float* positions_dev, float* out_dev, int64_t n, void* stream);

/* ------------------------------------------------------------------------------------------
 * Peer frame buffers (B200 addition, no reference analogue: the reference renders on one GPU). One process per GPU; the collecting
 * rank creates a device buffer and passes the 64-byte handle to the others (any byte transport: torch.distributed, a pipe); they
 * open it and pass the mapped address (plus their view / band offset) as `out_dev` of fvsrn_render_dvr: the DVR kernel's pixel
 * writes travel over NVLink into the collecting GPU's frame while the march is still running - render and gather are one
 * kernel per rank, no staging image and no separate transfer. Needs peer access between the GPUs (NVSwitch / NVLink).
 * The collecting rank reads the frame after a stream-ordered signal from every writer (fv-srn_b200/dist.py PeerFrames uses a
 * 4-byte NCCL all-reduce). create / free on the owner, open / close on the peers.
 * ------------------------------------------------------------------------------------------ */
int  fvsrn_peer_buffer_create(size_t bytes, void** dev_ptr, unsigned char handle[64]);
int  fvsrn_peer_buffer_open(const unsigned char handle[64], void** dev_ptr);
int  fvsrn_peer_buffer_close(void* dev_ptr);
int  fvsrn_peer_buffer_free(void* dev_ptr);

/* Launch diagnostics (stand-in for compute_max_warps / KernelLoader occupancy query) */
typedef struct fvsrn_query_info {
    int device, sm_count;
    int block_threads, blocks_per_sm, grid_blocks;
    int smem_bytes_per_block, regs_per_thread;
    int rays_per_warp_tile_x, rays_per_warp_tile_y;
    unsigned long long kernel_launches; /* number of fvsrn kernels launched by this handle so far */
} fvsrn_query_info;
int  fvsrn_query(fvsrn_network* net, int kernel /*0 = dvr, 1 = evaluate*/, fvsrn_query_info* info);

#ifdef __cplusplus
}
#endif
#endif /* FVSRN_B200_H */
