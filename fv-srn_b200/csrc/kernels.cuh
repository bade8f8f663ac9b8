// The two kernels of the hot path:
//   dvr_kernel   <- ImageEvaluatorSimpleKernel + RayEvaluationSteppingDvr::eval + TF + Blending
//                   (reference renderer_image_evaluator_simple.cuh:36-127, renderer_ray_evaluation_stepping_dvr.cuh:49-157)
//   eval_kernel  <- EvaluateNoBatches (reference renderer_volume_kernels1.cuh:15-71)
// Persistent CTAs; every warp owns an 8x4 pixel tile (32 rays) and marches it in lock-step, one SRN sample per
// lane per step = one 32-row MMA tile per step. Warps never synchronise with each other after the weights have
// been pinned in shared memory, so latency is hidden by warp-level parallelism.
#pragma once
#include "srn_core.cuh"
#include <cfloat>

namespace fvsrn {

constexpr int kWarpsPerBlock = 4;
constexpr int kTileW = 8, kTileH = 4;
constexpr int kTfSmemFloats = 1024; // 256 x float4 texture TF or up to 204 piecewise control points

struct DvrParams {
    int W, H;                 // full image
    int x0, y0, tw, th;       // rendered sub-rectangle
    int b0, nb;               // first batch index and number of batches of this launch
    const float* cam;         // (B,3,3)
    float fov_y, aspect;       // fovY in radians; tan(fovY/2) is taken on the device as the reference does
    float stepsize, alpha_early_out, dmin, inv_range, dmax;
    int early_out, blend_mode;
    int tf_kind, tf_R, tf_batch;
    const float* tf;
    int spp, spp_side;
    float* out;
    long long sb, sc, sy, sx;
    int* step_count;                    // optional (B,H,W)
    unsigned long long* counters;       // optional [valid, executed]
    unsigned int* tile_counter;         // work queue
    int tiles_x, tiles_y, num_tiles;
    int tf_smem_floats;                 // SS-mode tcgen05 kernel: floats reserved for the staged TF (multiple of 4)
    // fused colour extraction (reference IImageEvaluator::ExtractColor, iimage_evaluator.cpp:26-135, run as a second pass there)
    int out_format;                     // 0 (B,8,H,W) fp32 raw | 1 (B,4,H,W) fp32 | 2 (B,H,W) packed RGBA8 | 3 (B,H,W) half4
    int channel_mode;                   // IImageEvaluator::ChannelMode: 0 mask, 1 normal, 3 colour (2 = depth is not fusable)
    int tonemap;                        // ChannelColor only: ACES filmic curve + gamma (TonemappingKernel)
    float max_exposure;
    // normals and shading. normals != 0: the launch evaluates normals - a network that predicts them
    // (densitygrad* / densitycurvature* output modes), finite differences (NetParams::gradient_mode) or a BRDF that needs them
    int normals;
    // BRDFLambert, renderer_brdf_lambert.cuh:19-38 (scalar parameters): bit 0 gradient-magnitude scaling, bit 1 Phong
    int brdf_flags;
    int light_type;                     // BRDF_LIGHT_POINT 0 / BRDF_LIGHT_DIRECTION 1
    float light[3];                     // position or direction
    float brdf_mag_scaling, brdf_ambient, brdf_specular, brdf_mag_center, brdf_mag_radius;
    int brdf_spec_exp;
};

// ground-truth density grid (kernels in kernels_exact.cuh)
struct GridVolumeParams {
    unsigned long long tex;     // 1-channel 3-D texture: unnormalised coordinates, clamp, linear (volume.cpp:149-160)
    int rx, ry, rz;             // resolution
    float box_min[3], box_size[3];
};

struct EvalParams {
    const float* pos;   // (N,3)
    const float* dir;   // (N,3) or null
    float* out;         // (N,cout)
    long long n;
    int cout;
};

// copy the constant blob to shared memory (reference: VolumeInterpolationTensorcores ctor :401-564)
__device__ __forceinline__ void load_blob(const NetParams& P, uint4* sblob) {
    const int n16 = P.blob_bytes / 16;
    for (int i = threadIdx.x; i < n16; i += blockDim.x) sblob[i] = __ldg(P.blob + i);
}

__device__ __forceinline__ float4 tf_piecewise(const float* tf, int R, float d) { // reference renderer_tf_piecewise.cuh:27-52
    int i;
    for (i = 0; i < R - 2; ++i)
        if (tf[(i + 1) * 5 + 4] > d) break;
    const float* v0 = tf + i * 5;
    const float* v1 = v0 + 5;
    const float p0 = v0[4], p1 = v1[4];
    d = fminf(fmaxf(d, p0), p1);
    const float f = __fdividef(d - p0, p1 - p0);
    return make_float4(v0[0] + f * (v1[0] - v0[0]), v0[1] + f * (v1[1] - v0[1]), v0[2] + f * (v1[2] - v0[2]),
                       v0[3] + f * (v1[3] - v0[3]));
}
__device__ __forceinline__ float4 tf_texture(const float* tf, int R, float d) { // reference renderer_tf_texture.cuh:46-56
    const float x = d * float(R) - 0.5f;
    const float fl = floorf(x);
    const int di = int(fl);
    const float df = x - fl;
    const float* v0 = tf + 4 * min(max(di, 0), R - 1);
    const float* v1 = tf + 4 * min(max(di + 1, 0), R - 1);
    return make_float4(v0[0] + df * (v1[0] - v0[0]), v0[1] + df * (v1[1] - v0[1]), v0[2] + df * (v1[2] - v0[2]),
                       v0[3] + df * (v1[3] - v0[3]));
}

struct Ray { float3 eye, dir; float tmin, tmax; };
struct CamFrame { float3 eye, right, up, front; float tan_x, tan_y; };

// fp32 operations exactly as the reference kernel executes them. The reference is compiled with --use_fast_math
// (kernel_loader.cpp:229-241): flush-to-zero arithmetic, div.approx / rcp.approx / rsqrt.approx / sin.approx /
// cos.approx / ex2.approx. Spelled as PTX here so that this file's results do not depend on its own compile flags.
// Ray setup decides per ray how many steps fall inside the box (t <= tmax); with these operations in the
// reference's order that count is bit-identical to the reference kernel's for every ray.
namespace rf {
#define FV_RF2(name, ptx) __device__ __forceinline__ float name(float a, float b) { float r; asm(ptx " %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b)); return r; }
#define FV_RF1(name, ptx) __device__ __forceinline__ float name(float a) { float r; asm(ptx " %0, %1;" : "=f"(r) : "f"(a)); return r; }
FV_RF2(add, "add.ftz.f32") FV_RF2(sub, "sub.ftz.f32") FV_RF2(mul, "mul.ftz.f32") FV_RF2(div, "div.approx.ftz.f32")
FV_RF2(fmin, "min.ftz.f32") FV_RF2(fmax, "max.ftz.f32")
FV_RF1(rcp, "rcp.approx.ftz.f32") FV_RF1(rsqrt, "rsqrt.approx.ftz.f32") FV_RF1(sin, "sin.approx.ftz.f32")
FV_RF1(cos, "cos.approx.ftz.f32") FV_RF1(ex2, "ex2.approx.ftz.f32") FV_RF1(floor, "cvt.rmi.ftz.f32.f32")
FV_RF1(lg2, "lg2.approx.ftz.f32") FV_RF1(neg, "neg.ftz.f32") FV_RF1(abs, "abs.ftz.f32")
__device__ __forceinline__ float fma(float a, float b, float c) { float r; asm("fma.rn.ftz.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c)); return r; }
// (a < b) ? a : b and (a > b) ? a : b as the reference's fminf/fmaxf-free helper_math compiles (setp + selp)
__device__ __forceinline__ float sel_lt(float a, float b) { return (a < b) ? a : b; }
__device__ __forceinline__ float sel_gt(float a, float b) { return (a > b) ? a : b; }
#undef FV_RF1
#undef FV_RF2
} // namespace rf

// camera frame, reference renderer_camera.cuh:34-52 (operation order of the compiled reference kernel)
__device__ __forceinline__ CamFrame load_camera(const float* cam, float fov_y, float aspect, int b) {
    CamFrame c;
    const float* cm = cam + (size_t)b * 9;
    c.eye = make_float3(__ldg(cm + 0), __ldg(cm + 1), __ldg(cm + 2));
    c.right = make_float3(__ldg(cm + 3), __ldg(cm + 4), __ldg(cm + 5));
    c.up = make_float3(__ldg(cm + 6), __ldg(cm + 7), __ldg(cm + 8));
    c.front = make_float3(rf::sub(rf::mul(c.right.z, c.up.y), rf::mul(c.right.y, c.up.z)),
                          rf::sub(rf::mul(c.right.x, c.up.z), rf::mul(c.right.z, c.up.x)),
                          rf::sub(rf::mul(c.right.y, c.up.x), rf::mul(c.right.x, c.up.y))); // cross(up, right)
    const float half_fov = rf::div(fov_y, 2.0f);
    c.tan_y = rf::div(rf::sin(half_fov), rf::cos(half_fov));      // tanf under fast math
    c.tan_x = rf::mul(c.tan_y, aspect);
    return c;
}
__device__ __forceinline__ CamFrame load_camera(const DvrParams& D, int b) { return load_camera(D.cam, D.fov_y, D.aspect, b); }

// direction of the camera ray through pixel (x,y), sub-sample s (reference renderer_image_evaluator_simple.cuh:84-86 =
// renderer_camera_kernels.cuh:23-30, renderer_camera.cuh:41-51); the origin is c.eye
__device__ __forceinline__ float3 camera_ray(int W, int H, const CamFrame& c, int x, int y, int s, int spp, int spp_side) {
    float nx = float(x), ny = float(y);
    if (spp > 1) {                // deterministic stratified sub-pixel offsets (no reference counterpart)
        nx = __fadd_rn(nx, __fsub_rn(__fdiv_rn(float(s % spp_side) + 0.5f, float(spp_side)), 0.5f));
        ny = __fadd_rn(ny, __fsub_rn(__fdiv_rn(float(s / spp_side) + 0.5f, float(spp_side)), 0.5f));
    }
    const float hx = rf::add(nx, 0.5f), hy = rf::add(ny, 0.5f);
    const float ndc_x = rf::add(rf::div(rf::add(hx, hx), float(W)), -1.0f);
    const float ndc_y = rf::add(rf::div(rf::add(hy, hy), float(H)), -1.0f);
    const float sx = rf::mul(ndc_x, c.tan_x), sy = rf::mul(ndc_y, c.tan_y);
    float3 dir = make_float3(rf::fma(sy, c.up.x, rf::fma(sx, c.right.x, c.front.x)),
                             rf::fma(sy, c.up.y, rf::fma(sx, c.right.y, c.front.y)),
                             rf::fma(sy, c.up.z, rf::fma(sx, c.right.z, c.front.z)));
    const float len2 = rf::fma(dir.z, dir.z, rf::fma(dir.x, dir.x, rf::mul(dir.y, dir.y)));
    const float inv_len = rf::rsqrt(len2);
    dir.x = rf::mul(inv_len, dir.x); dir.y = rf::mul(inv_len, dir.y); dir.z = rf::mul(inv_len, dir.z);
    return dir;
}
// the ray and its box interval (renderer_utils.cuh:91-105, renderer_ray_evaluation_stepping_dvr.cuh:66-71)
__device__ __forceinline__ Ray make_ray(const DvrParams& D, const CamFrame& c, int x, int y, int s, float3 bmin, float3 bsize, int spp) {
    float3 dir = camera_ray(D.W, D.H, c, x, y, s, spp, D.spp_side);
    const float3 eye = c.eye;
    const float ix = rf::div(1.0f, dir.x), iy = rf::div(1.0f, dir.y), iz = rf::div(1.0f, dir.z);
    const float t1 = rf::mul(rf::sub(bmin.x, eye.x), ix), t2 = rf::mul(ix, rf::sub(rf::add(bmin.x, bsize.x), eye.x));
    const float t3 = rf::mul(rf::sub(bmin.y, eye.y), iy), t4 = rf::mul(iy, rf::sub(rf::add(bmin.y, bsize.y), eye.y));
    const float t5 = rf::mul(rf::sub(bmin.z, eye.z), iz), t6 = rf::mul(rf::sub(rf::add(bmin.z, bsize.z), eye.z), iz);
    Ray r; r.eye = eye; r.dir = dir;
    r.tmin = rf::fmax(rf::sel_gt(rf::sel_gt(rf::sel_lt(t1, t2), rf::sel_lt(t3, t4)), rf::sel_lt(t5, t6)), 0.f);
    r.tmax = rf::fmin(rf::sel_lt(rf::sel_lt(rf::sel_gt(t1, t2), rf::sel_gt(t3, t4)), rf::sel_gt(t5, t6)), FLT_MAX);
    return r;
}

// density of a density-type network from its raw first output (renderer_volume_tensorcores.cuh:1073-1158): the plain modes
// squash it with a sigmoid, the ":direct" / ":cubic" modes pass it on
__device__ __forceinline__ bool output_mode_is_colour(int m) { return m == D_OUT_RGBO || m == D_OUT_RGBO_DIRECT; }
__device__ __forceinline__ bool output_mode_has_sigmoid(int m) { return m == D_OUT_DENSITY || m == D_OUT_DENSITY_GRADIENT || m == D_OUT_DENSITY_CURVATURE; }
__device__ __forceinline__ float density_from_output(int mode, float o0) { return output_mode_has_sigmoid(mode) ? sigmoidf_(o0) : o0; }

// network output -> colour + absorption of one sample (TF or colour network), reference
// renderer_ray_evaluation_stepping_dvr.cuh:100-140, renderer_volume_tensorcores.cuh:1054-1072, 1149-1158
__device__ __forceinline__ float4 shade_sample(const NetParams& P, const DvrParams& D, const float* tfp, const float (&o)[4]) {
    float4 c = make_float4(0.f, 0.f, 0.f, 0.f);
    if (!output_mode_is_colour(P.output_mode)) {
        const float value = density_from_output(P.output_mode, o[0]);
        if (value >= D.dmin) {
            const float d2 = fminf(fmaxf((value - D.dmin) * D.inv_range, 0.f), 1.f);
            if (D.tf_kind == 0) c = tf_piecewise(tfp, D.tf_R, d2);
            else if (D.tf_kind == 1) c = tf_texture(tfp, D.tf_R, d2);
            else c = make_float4(d2 * tfp[1], d2 * tfp[1], d2 * tfp[1], d2 * tfp[0]);
            c.w *= D.stepsize;
        }
    } else {    // colour network, SKIP_TRANSFER_FUNCTION path
        if (P.output_mode == D_OUT_RGBO) c = make_float4(sigmoidf_(o[0]), sigmoidf_(o[1]), sigmoidf_(o[2]), softplusf_(o[3]));
        else c = make_float4(fminf(fmaxf(o[0], 0.f), 1.f), fminf(fmaxf(o[1], 0.f), 1.f), fminf(fmaxf(o[2], 0.f), 1.f), fmaxf(o[3], 0.f));
        c.w *= D.stepsize;
    }
    return c;
}
// front-to-back compositing, reference renderer_blending.cuh:36-51 (normal is zero for density networks)
__device__ __forceinline__ void blend_sample(const DvrParams& D, float4 c, float t, float4& acc, float& acc_depth) {
    if (c.w > 0.f) {
        const float a = (D.blend_mode == 1) ? 1.f - __expf(-c.w) : fminf(1.f, c.w);
        const float k = (1.f - acc.w) * a;
        acc.x += k * c.x; acc.y += k * c.y; acc.z += k * c.z;
        acc_depth += k * t;
        acc.w += k;
    }
}
// ---- normals and shading
// normal predicted by the network next to the density (:1077-1127); densitygrad:cubic stores the cube root
__device__ __forceinline__ float3 normal_from_output(int mode, const float (&o)[4]) {
    if (mode == D_OUT_DENSITY_GRADIENT_CUBIC) return make_float3(o[1] * o[1] * o[1], o[2] * o[2] * o[2], o[3] * o[3] * o[3]);
    return make_float3(o[1], o[2], o[3]);
}
// safeNormalize, helper_math.cuh:2443-2448
__device__ __forceinline__ float3 safe_normalize(float3 v) {
    const float l2 = v.x * v.x + v.y * v.y + v.z * v.z;
    if (l2 < 1e-8f) return v;
    const float inv = rsqrtf(l2);
    return make_float3(v.x * inv, v.y * inv, v.z * inv);
}
// BRDFLambert::eval, renderer_brdf_lambert.cuh:57-103: gradient-magnitude scaling of the absorption and Phong shading of
// the colour. `gradientNorm` is rsqrt(|g|^2) there (:63) and enters the smoothstep as such - kept.
__device__ __forceinline__ float4 brdf_lambert(const DvrParams& D, float4 c, float3 wpos, float3 g, float3 dir) {
    const float g2 = g.x * g.x + g.y * g.y + g.z * g.z;
    const float ginv = rsqrtf(g2);
    const float3 n = safe_normalize(g);
    if (D.brdf_flags & 1) c.w *= 1.f - __expf(-D.brdf_mag_scaling * g2);
    if (D.brdf_flags & 2) {
        float3 l = (D.light_type == 1) ? make_float3(-D.light[0], -D.light[1], -D.light[2])
                                       : make_float3(D.light[0] - wpos.x, D.light[1] - wpos.y, D.light[2] - wpos.z);
        const float linv = rsqrtf(l.x * l.x + l.y * l.y + l.z * l.z);
        l.x *= linv; l.y *= linv; l.z *= linv;
        const float lo = D.brdf_mag_center - D.brdf_mag_radius, hi = D.brdf_mag_center + D.brdf_mag_radius;
        const float y = fminf(fmaxf(__fdividef(ginv - lo, hi - lo), 0.f), 1.f);
        const float phong = y * y * (3.f - 2.f * y);                             // smoothstep
        const float amb = 1.f + phong * (D.brdf_ambient - 1.f);                  // lerp(1, ambient, phong)
        const float ndl = n.x * l.x + n.y * l.y + n.z * l.z;
        // reflect(l, -n) = l - 2 (-n) dot(-n, l)
        const float k = 2.f * ndl;
        const float3 r = make_float3(l.x - k * n.x, l.y - k * n.y, l.z - k * n.z);
        const float rd = fmaxf(0.f, dir.x * r.x + dir.y * r.y + dir.z * r.z);
        const float e = float(D.brdf_spec_exp);
        const float spec = D.brdf_specular * float(double(e + 2.f) * 0.159155) * __powf(rd, e);
        const float d = fabsf(ndl);
        c.x = amb * c.x + (1.f - amb) * (d * c.x + spec);
        c.y = amb * c.y + (1.f - amb) * (d * c.y + spec);
        c.z = amb * c.z + (1.f - amb) * (d * c.z + spec);
    }
    return c;
}
// Blending::eval with the normal channel, renderer_blending.cuh:36-51
__device__ __forceinline__ void blend_sample_normal(const DvrParams& D, float4 c, float3 n, float t, float4& acc, float3& acc_n, float& acc_depth) {
    const float a = (D.blend_mode == 1) ? 1.f - __expf(-c.w) : fminf(1.f, c.w);
    const float k = (1.f - acc.w) * a;
    acc.x += k * c.x; acc.y += k * c.y; acc.z += k * c.z;
    acc_n.x += k * n.x; acc_n.y += k * n.y; acc_n.z += k * n.z;
    acc_depth += k * t;
    acc.w += k;
}

// the same compositing step without divergent branches: `on` = this sample exists (valid step of a live ray)
__device__ __forceinline__ void blend_sample_pred(const DvrParams& D, int blend_mode, bool on, float4 c, float t, float4& acc, float& acc_depth) {
    float e;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(c.w * -1.4426950408889634f));
    const float a = (blend_mode == 1) ? 1.f - e : fminf(1.f, c.w);
    const float k = (on && c.w > 0.f) ? (1.f - acc.w) * a : 0.f;
    acc.x = fmaf(k, c.x, acc.x); acc.y = fmaf(k, c.y, acc.y); acc.z = fmaf(k, c.z, acc.z);
    acc_depth = fmaf(k, t, acc_depth);
    acc.w += k;
}
// ACES filmic curve + gamma, reference tonemappingFunction (iimage_evaluator_cuda.cu:145-166; that file is built with
// --use_fast_math, CMakeLists.txt:44, hence the approximate division and power)
__device__ __forceinline__ float tonemap1(float v, float max_exposure) {
    v = rf::div(v, max_exposure);
    v = rf::div(v * (2.51f * v + 0.03f), v * (2.43f * v + 0.59f) + 0.14f);
    v = fminf(fmaxf(v, 0.f), 1.f);
    float l; asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(l) : "f"(v));
    return rf::ex2(l * (1.0f / 2.4f));
}
// rgbaToInt, renderer_utils.cuh:48-57
__device__ __forceinline__ unsigned int rgba_to_int(float r, float g, float b, float a) {
    r = fminf(fmaxf(r * 255.f, 0.f), 255.f); g = fminf(fmaxf(g * 255.f, 0.f), 255.f);
    b = fminf(fmaxf(b * 255.f, 0.f), 255.f); a = fminf(fmaxf(a * 255.f, 0.f), 255.f);
    return ((unsigned int)a << 24) | ((unsigned int)b << 16) | ((unsigned int)g << 8) | (unsigned int)r;
}
// one pixel of the selected channel: what SelectOutputChannelKernel / TonemappingKernel (iimage_evaluator_cuda.cu:21-80,
// 168-260) compute from the 8 raw channels, with the channel table of IImageEvaluator::ExtractColor (iimage_evaluator.cpp:52-112)
__device__ __forceinline__ float4 select_channel(int channel_mode, int tonemap, float max_exposure, float4 color, float3 normal) {
    if (channel_mode == 3) {        // ChannelColor
        if (tonemap) return make_float4(tonemap1(color.x, max_exposure), tonemap1(color.y, max_exposure), tonemap1(color.z, max_exposure), color.w);
        return color;
    }
    if (channel_mode == 0) return make_float4(color.w, color.w, color.w, 1.f);                                   // ChannelMask
    return make_float4(fmaf(normal.x, 0.5f, 0.5f), fmaf(normal.y, 0.5f, 0.5f), fmaf(normal.z, 0.5f, 0.5f), color.w);  // ChannelNormal
}
__device__ __forceinline__ void store_rgba(const DvrParams& D, int b, int x, int y, float4 c) {
    const long long o = (long long)b * D.sb + (long long)y * D.sy + (long long)x * D.sx;
    if (D.out_format == 1) {
        float* p = D.out + o;
        p[0 * D.sc] = c.x; p[1 * D.sc] = c.y; p[2 * D.sc] = c.z; p[3 * D.sc] = c.w;
    } else if (D.out_format == 2) {
        reinterpret_cast<unsigned int*>(D.out)[o] = rgba_to_int(c.x, c.y, c.z, c.w);
    } else {
        const __half2 lo = __floats2half2_rn(c.x, c.y), hi = __floats2half2_rn(c.z, c.w);
        reinterpret_cast<uint2*>(D.out)[o] = make_uint2(h2u(lo), h2u(hi));
    }
}
// pixel write, reference renderer_image_evaluator_simple.cuh:101-124 (+ the fused colour extraction when requested)
__device__ __forceinline__ void write_pixel(const DvrParams& D, int b, int x, int y, float4 pix_color, float pix_depth, int last_steps, int spp,
                                            float3 pix_normal = make_float3(0.f, 0.f, 0.f)) {
    const float inv_s = 1.f / float(spp);
    if (D.step_count) D.step_count[((size_t)b * D.H + y) * D.W + x] = last_steps;
    const float4 color = make_float4(pix_color.x * inv_s, pix_color.y * inv_s, pix_color.z * inv_s, pix_color.w * inv_s);
    const float3 normal = make_float3(pix_normal.x * inv_s, pix_normal.y * inv_s, pix_normal.z * inv_s);
    if (D.out_format != 0) {
        store_rgba(D, b, x, y, select_channel(D.channel_mode, D.tonemap, D.max_exposure, color, normal));
        return;
    }
    float* o = D.out + (long long)b * D.sb + (long long)y * D.sy + (long long)x * D.sx;
    const float depth = pix_depth / pix_color.w;    // NaN for empty pixels, as in the reference
    o[0 * D.sc] = color.x; o[1 * D.sc] = color.y; o[2 * D.sc] = color.z;
    o[3 * D.sc] = color.w;
    o[4 * D.sc] = normal.x; o[5 * D.sc] = normal.y; o[6 * D.sc] = normal.z;
    o[7 * D.sc] = depth;
}

// NORMALS: the launch evaluates normals (predicted by the network or finite differences, BRDFLambert shading,
// normal channels of the output) - its own instantiation, the plain march keeps its registers
template <int H16, int GC16, bool DIR, bool NORMALS>
__global__ void __launch_bounds__(kWarpsPerBlock * 32, NORMALS ? 2 : 4)
dvr_kernel(const __grid_constant__ NetParams P, const __grid_constant__ DvrParams D) {
    using Net = Srn<H16, GC16, DIR>;
    extern __shared__ uint4 smem[];
    uint4* sblob = smem;
    const int blob16 = (P.blob_bytes + 15) / 16;
    float* stf = reinterpret_cast<float*>(smem + blob16);
    uint4* stage_all = smem + blob16 + kTfSmemFloats / 4;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint4* stage = stage_all + warp * (Net::STAGE_BYTES / 16);
    // adjoint gradient mode only (the launch adds the bytes): this warp's pre-activation store
    uint32_t* pre = reinterpret_cast<uint32_t*>(stage_all + kWarpsPerBlock * (Net::STAGE_BYTES / 16)) + warp * Net::pre_words(P.num_hidden);

    load_blob(P, sblob);
    const int tf_floats = D.tf_kind == 0 ? D.tf_R * 5 : (D.tf_kind == 1 ? D.tf_R * 4 : 2);
    if (D.tf_batch == 1)
        for (int i = threadIdx.x; i < tf_floats; i += blockDim.x) stf[i] = __ldg(D.tf + i);
    __syncthreads();

    const uint32_t sblob_addr = smem_u32(sblob);
    const ActConst ac = make_act_const(P.activation, P.act_param);
    const float3 bmin = make_float3(P.box_min[0], P.box_min[1], P.box_min[2]);
    const float3 bsize = make_float3(P.box_size[0], P.box_size[1], P.box_size[2]);
    const float3 binv = make_float3(__fdiv_rn(1.f, bsize.x), __fdiv_rn(1.f, bsize.y), __fdiv_rn(1.f, bsize.z));
    unsigned long long n_valid = 0, n_exec = 0;
    for (;;) {
        int tile = 0;
        if (lane == 0) tile = int(atomicAdd(D.tile_counter, 1u));
        tile = __shfl_sync(0xffffffffu, tile, 0);
        if (tile >= D.num_tiles) break;
        const int tiles_per_img = D.tiles_x * D.tiles_y;
        const int bl = tile / tiles_per_img;            // local batch
        const int trem = tile - bl * tiles_per_img;
        const int ty = trem / D.tiles_x, tx = trem - ty * D.tiles_x;
        const int b = D.b0 + bl;
        const int x = D.x0 + tx * kTileW + (lane & (kTileW - 1));
        const int y = D.y0 + ty * kTileH + (lane / kTileW);
        const bool in_img = (x < D.x0 + D.tw) && (y < D.y0 + D.th);
        const float* tfp = (D.tf_batch == 1) ? stf : (D.tf + (size_t)b * tf_floats);

        const CamFrame cf = load_camera(D, b);

        float4 pix_color = make_float4(0.f, 0.f, 0.f, 0.f);
        float3 pix_normal = make_float3(0.f, 0.f, 0.f);
        float pix_depth = 0.f;
        int last_steps = 0;
        for (int s = 0; s < D.spp; ++s) {   // warp-uniform trip count
            const Ray r = make_ray(D, cf, x, y, s, bmin, bsize, D.spp);
            const float3 eye = r.eye, dir = r.dir;
            const float tmin = r.tmin;
            const float tmax = in_img ? r.tmax : -1.f;  // lanes outside the image never become valid

            float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
            float3 acc_n = make_float3(0.f, 0.f, 0.f);
            float acc_depth = 0.f;
            int steps = 0;
            // lock-step march, reference renderer_ray_evaluation_stepping_dvr.cuh:84-154
            for (int i = 0;; ++i) {
                const float t = __fmaf_rn(float(i), D.stepsize, tmin);
                bool valid = (t <= tmax);
                if (D.early_out) valid = valid && (acc.w < D.alpha_early_out);
                if (!__any_sync(0xffffffffu, valid)) break;
                n_exec += 32;
                float px = 0.f, py = 0.f, pz = 0.f;
                if (valid) {
                    px = __fmul_rn(__fsub_rn(__fmaf_rn(dir.x, t, eye.x), bmin.x), binv.x);
                    py = __fmul_rn(__fsub_rn(__fmaf_rn(dir.y, t, eye.y), bmin.y), binv.y);
                    pz = __fmul_rn(__fsub_rn(__fmaf_rn(dir.z, t, eye.z), bmin.z), binv.z);
                }
                float o[4];
                Net::forward(P, sblob_addr, stage, ac, px, py, pz, dir.x, dir.y, dir.z, o);
                if constexpr (!NORMALS) {
                    if (!valid) continue;
                    ++steps;
                    blend_sample(D, shade_sample(P, D, tfp, o), t, acc, acc_depth);
                } else {
                // normals, :114-153: the normal is evaluated when ANY lane of the warp shades its sample
                const float value = density_from_output(P.output_mode, o[0]);
                const bool shaded = valid && value >= D.dmin;
                float3 n = make_float3(0.f, 0.f, 0.f);
                if (P.gradient_mode == 1) {     // GRADIENT_MODE_FINITE_DIFFERENCES, renderer_volume_tensorcores.cuh:1185-1197
                    if (__any_sync(0xffffffffu, shaded)) {
                        const float h = P.fd_stepsize;
                        float d[6];
#pragma unroll 1
                        for (int k = 0; k < 6; ++k) {
                            const float sgn = (k & 1) ? -h : h;
                            const float wx = __fmaf_rn(dir.x, t, eye.x) + ((k >> 1) == 0 ? sgn : 0.f);
                            const float wy = __fmaf_rn(dir.y, t, eye.y) + ((k >> 1) == 1 ? sgn : 0.f);
                            const float wz = __fmaf_rn(dir.z, t, eye.z) + ((k >> 1) == 2 ? sgn : 0.f);
                            float q[4];
                            Net::forward(P, sblob_addr, stage, ac, valid ? (wx - bmin.x) * binv.x : 0.f, valid ? (wy - bmin.y) * binv.y : 0.f,
                                         valid ? (wz - bmin.z) * binv.z : 0.f, dir.x, dir.y, dir.z, q);
                            const float v = density_from_output(P.output_mode, q[0]);
                            if (k == 0) d[0] = v; else if (k == 1) d[1] = v; else if (k == 2) d[2] = v;
                            else if (k == 3) d[3] = v; else if (k == 4) d[4] = v; else d[5] = v;
                        }
                        const float inv2h = __fdividef(1.f, 2.f * h);
                        n = make_float3((d[0] - d[1]) * inv2h, (d[2] - d[3]) * inv2h, (d[4] - d[5]) * inv2h);
                    }
                } else if (!DIR && GC16 > 0 && P.gradient_mode == 2) {
                    // GRADIENT_MODE_ADJOINT_METHOD (:1199-1530): on steps that shade something the forward pass is repeated with its
                    // pre-activations kept, then one backward pass gives d density / d position (unit-box coordinates, as the reference)
                    if constexpr (!DIR && GC16 > 0) {
                        if (__any_sync(0xffffffffu, shaded)) {
                            float q[4];
                            Net::template forward<true>(P, sblob_addr, stage, ac, px, py, pz, dir.x, dir.y, dir.z, q, pre);
                            n = Net::adjoint(P, sblob_addr, stage, pre, ac, px, py, pz, q[0]);
                        }
                    }
                } else if (P.output_mode >= D_OUT_DENSITY_GRADIENT) n = normal_from_output(P.output_mode, o);
                if (valid) ++steps;
                if (!shaded) continue;
                float4 c = shade_sample(P, D, tfp, o);
                if (c.w > 0.f) {
                    if (D.brdf_flags) c = brdf_lambert(D, c, make_float3(__fmaf_rn(dir.x, t, eye.x), __fmaf_rn(dir.y, t, eye.y), __fmaf_rn(dir.z, t, eye.z)), n, dir);
                    blend_sample_normal(D, c, safe_normalize(n), t, acc, acc_n, acc_depth);
                }
                }
            }
            n_valid += steps;
            last_steps = steps;
            pix_color.x += acc.x; pix_color.y += acc.y; pix_color.z += acc.z; pix_color.w += acc.w;
            pix_normal.x += acc_n.x * acc.w; pix_normal.y += acc_n.y * acc.w; pix_normal.z += acc_n.z * acc.w;
            pix_depth += acc_depth * acc.w;     // reference renderer_image_evaluator_simple.cuh:101-103
        }
        if (in_img) write_pixel(D, b, x, y, pix_color, pix_depth, last_steps, D.spp, pix_normal);
    }
    if (D.counters) {
        for (int off = 16; off > 0; off >>= 1) n_valid += __shfl_xor_sync(0xffffffffu, n_valid, off);
        if (lane == 0) { atomicAdd(D.counters + 0, n_valid); atomicAdd(D.counters + 1, n_exec); }
    }
}

template <int H16, int GC16, bool DIR>
__global__ void __launch_bounds__(kWarpsPerBlock * 32)
eval_kernel(const __grid_constant__ NetParams P, const __grid_constant__ EvalParams E) {
    using Net = Srn<H16, GC16, DIR>;
    extern __shared__ uint4 smem[];
    uint4* sblob = smem;
    const int blob16 = (P.blob_bytes + 15) / 16;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint4* stage = smem + blob16 + warp * (Net::STAGE_BYTES / 16);
    load_blob(P, sblob);
    __syncthreads();
    const uint32_t sblob_addr = smem_u32(sblob);
    const ActConst ac = make_act_const(P.activation, P.act_param);
    const long long warps_total = (long long)gridDim.x * kWarpsPerBlock;
    const long long n32 = (E.n + 31) / 32;
    for (long long w = (long long)blockIdx.x * kWarpsPerBlock + warp; w < n32; w += warps_total) {
        const long long i = w * 32 + lane;
        const bool valid = i < E.n;
        float px = 0.f, py = 0.f, pz = 0.f, dx = 0.f, dy = 0.f, dz = 0.f;
        if (valid) {
            px = __ldg(E.pos + 3 * i); py = __ldg(E.pos + 3 * i + 1); pz = __ldg(E.pos + 3 * i + 2);
            if (DIR && E.dir) { dx = __ldg(E.dir + 3 * i); dy = __ldg(E.dir + 3 * i + 1); dz = __ldg(E.dir + 3 * i + 2); }
        }
        float o[4];
        Net::forward(P, sblob_addr, stage, ac, px, py, pz, dx, dy, dz, o);
        if (!valid) continue;
        // output parametrisation, reference :1054-1072, 1149-1158 (value, not valueNoClamping)
        if (!output_mode_is_colour(P.output_mode)) E.out[i] = density_from_output(P.output_mode, o[0]);
        else {
            float4 c;
            if (P.output_mode == D_OUT_RGBO) c = make_float4(sigmoidf_(o[0]), sigmoidf_(o[1]), sigmoidf_(o[2]), softplusf_(o[3]));
            else c = make_float4(fminf(fmaxf(o[0], 0.f), 1.f), fminf(fmaxf(o[1], 0.f), 1.f), fminf(fmaxf(o[2], 0.f), 1.f), fmaxf(o[3], 0.f));
            reinterpret_cast<float4*>(E.out)[i] = c;
        }
    }
}

} // namespace fvsrn
