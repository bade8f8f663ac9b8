// AOT sm_100a instantiations of the reference-exact DVR kernel (kernels_exact.cuh). This unit is compiled with
// --use_fast_math, like the reference compiles its kernels (kernel_loader.cpp:229-241), so that the header-implemented
// half functions it shares with the reference (hsin, hcos, __hdiv of cuda_fp16.hpp) expand to the same instructions.
#include "launch_exact.cuh"
namespace fvsrn {

// ---- tensor operators beside the fused path, in the reference's fast-math operation order
// CameraGenerateRayKernel, renderer_camera_kernels.cuh:12-45: (B,H,W,3) ray origins and directions
__global__ void generate_rays_kernel(const float* __restrict__ cam, float fov_y, float aspect, int B, int H, int W,
                                     float* __restrict__ start, float* __restrict__ dir) {
    const long long total = (long long)B * H * W;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int x = int(i % W), y = int((i / W) % H), b = int(i / ((long long)W * H));
        const CamFrame c = load_camera(cam, fov_y, aspect, b);
        const float3 d = camera_ray(W, H, c, x, y, 0, 1, 1);
        start[3 * i + 0] = c.eye.x; start[3 * i + 1] = c.eye.y; start[3 * i + 2] = c.eye.z;
        dir[3 * i + 0] = d.x; dir[3 * i + 1] = d.y; dir[3 * i + 2] = d.z;
    }
}
// EvaluateTF / EvaluateTFWithPrevious, renderer_tf_kernels.cuh:11-70 (batch 0 of the TF; the previous density only matters
// for pre-integrated texture TFs, which are not on the path): densities (N) -> colours (N,4), absorption times stepsize
__global__ void evaluate_tf_kernel(int tf_kind, const float* __restrict__ tf, int R, const float* __restrict__ density, float* __restrict__ out,
                                   long long n, float dmin, float dmax, float stepsize) {
    const float inv_range = rf::rcp(rf::sub(dmax, dmin));
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const float d = density[i];
        float4 c = make_float4(0.f, 0.f, 0.f, 0.f);
        if (d >= dmin) {
            const float d2 = rf::fmax(0.f, rf::fmin(rf::mul(rf::sub(d, dmin), inv_range), 1.0f));
            if (tf_kind == 0) c = exact_tf_piecewise(tf, R, d2, stepsize);
            else if (tf_kind == 1) c = exact_tf_texture(tf, R, d2, stepsize);
            else c = make_float4(d2 * tf[1], d2 * tf[1], d2 * tf[1], d2 * tf[0] * stepsize);      // renderer_tf_identity.cuh
        }
        reinterpret_cast<float4*>(out)[i] = c;
    }
}
cudaError_t launch_generate_rays(const float* cam, float fov_y, float aspect, int B, int H, int W, float* start, float* dir, cudaStream_t s) {
    const long long total = (long long)B * H * W;
    if (total <= 0) return cudaSuccess;
    generate_rays_kernel<<<int(std::min<long long>((total + 255) / 256, 148 * 8)), 256, 0, s>>>(cam, fov_y, aspect, B, H, W, start, dir);
    return cudaGetLastError();
}
cudaError_t launch_evaluate_tf(int tf_kind, const float* tf, int R, const float* density, float* out, long long n, float dmin, float dmax,
                               float stepsize, cudaStream_t s) {
    if (n <= 0) return cudaSuccess;
    evaluate_tf_kernel<<<int(std::min<long long>((n + 255) / 256, 148 * 8)), 256, 0, s>>>(tf_kind, tf, R, density, out, n, dmin, dmax, stepsize);
    return cudaGetLastError();
}

cudaError_t launch_dvr_grid(const GridVolumeParams& G, const DvrParams& D, int sm_count, cudaStream_t s) {
    const long long total = (long long)D.tw * D.th * D.nb;
    if (total <= 0) return cudaSuccess;
    dvr_grid_kernel<<<int(std::min<long long>((total + 255) / 256, (long long)sm_count * 8)), 256, 0, s>>>(G, D);
    return cudaGetLastError();
}
cudaError_t launch_eval_grid(const GridVolumeParams& G, const float* pos, float* out, long long n, cudaStream_t s) {
    if (n <= 0) return cudaSuccess;
    eval_grid_kernel<<<int(std::min<long long>((n + 255) / 256, 148 * 16)), 256, 0, s>>>(G, pos, out, n);
    return cudaGetLastError();
}

cudaError_t launch_exact(int h16, int gc16, int dir, const NetParams& P, const DvrParams& D, LaunchCfg* cfg, cudaStream_t s) {
    if (gc16 != 1 || dir) {     // two latent groups / direction input: the 32- and 64-wide networks
        if (gc16 < 1 || gc16 > 2 || (h16 != 2 && h16 != 4)) return cudaErrorInvalidValue;
        if (h16 == 2) return gc16 == 2 ? (dir ? launch_exact_h<2, 2, true>(P, D, cfg, s) : launch_exact_h<2, 2, false>(P, D, cfg, s)) : launch_exact_h<2, 1, true>(P, D, cfg, s);
        return gc16 == 2 ? (dir ? launch_exact_h<4, 2, true>(P, D, cfg, s) : launch_exact_h<4, 2, false>(P, D, cfg, s)) : launch_exact_h<4, 1, true>(P, D, cfg, s);
    }
    switch (h16) {
        case 2: return launch_exact_h<2>(P, D, cfg, s);
        case 3: return launch_exact_h<3>(P, D, cfg, s);
        case 4: return launch_exact_h<4>(P, D, cfg, s);
        case 6: return launch_exact_h<6>(P, D, cfg, s);
        case 8: return launch_exact_h<8>(P, D, cfg, s);
    }
    return cudaErrorInvalidValue;
}
} // namespace fvsrn
