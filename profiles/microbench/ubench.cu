// Micro-benchmarks that decide the design of the tcgen05 SRN kernel (run on the B200 box, see profiles/README.md):
//   1. tcgen05.ld throughput (accumulator read-back in the per-layer epilogue)
//   2. latent fetch: hardware-filtered fp16 RGBA 3-D textures vs. manual 8-corner LDG.128 + HFMA2 blend
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ubench ubench.cu
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <vector>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); return 1; } } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// ---------------------------------------------------------------- 1. TMEM read / write throughput
template <int MODE>  // 0: ld x32 only, 1: ld x32 + st x16 (epilogue pattern)
__global__ void __launch_bounds__(128) tmem_kernel(int iters, unsigned long long* cycles, uint32_t* sink) {
    __shared__ uint32_t taddr_s;
    if (threadIdx.x < 32) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;\n" :: "r"(smem_u32(&taddr_s)), "r"(64u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory");
    const uint32_t taddr = taddr_s + ((uint32_t(threadIdx.x >> 5) * 32u) << 16);
    uint32_t acc = 0;
    uint32_t r[32];
    const long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 "
                     "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];\n"
                     : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                       "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
                       "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
                       "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
                     : "r"(taddr) : "memory");
        asm volatile("tcgen05.wait::ld.sync.aligned;\n" ::: "memory");
#pragma unroll
        for (int j = 0; j < 32; ++j) acc ^= r[j];
        if (MODE == 1) {
            asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};\n"
                         :: "r"(taddr + 32), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]),
                            "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]) : "memory");
            asm volatile("tcgen05.wait::st.sync.aligned;\n" ::: "memory");
        }
    }
    const long long t1 = clock64();
    if (threadIdx.x == 0) cycles[blockIdx.x] = (unsigned long long)(t1 - t0);
    if (acc == 0x12345678u) sink[0] = acc;
    asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory");
    __syncthreads();
    if (threadIdx.x < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;\n" :: "r"(taddr_s), "r"(64u) : "memory");
}

// ---------------------------------------------------------------- 2. latent fetch
struct Tex4 { cudaTextureObject_t t[4]; };

__device__ __forceinline__ void ray_pos(int i, int step, float& x, float& y, float& z) {
    // coherent pseudo-rays: neighbouring threads start next to each other and walk the same direction
    const int px = i & 1023, py = (i >> 10) & 1023;
    x = px * (1.f / 1024.f) * 0.6f + step * (0.4f / 512.f);
    y = py * (1.f / 1024.f) * 0.7f + step * (0.3f / 512.f);
    z = step * (1.f / 512.f);
}

__global__ void __launch_bounds__(128) fetch_tex_kernel(Tex4 T, int steps, float* out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    __half2 acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = __float2half2_rn(0.f);
    for (int s = 0; s < steps; ++s) {
        float x, y, z; ray_pos(i, s, x, y, z);
#pragma unroll
        for (int g = 0; g < 4; ++g) {
            const float4 v = tex3D<float4>(T.t[g], x, y, z);
            acc[2 * g] = __hadd2(acc[2 * g], __floats2half2_rn(v.x, v.y));
            acc[2 * g + 1] = __hadd2(acc[2 * g + 1], __floats2half2_rn(v.z, v.w));
        }
    }
    float r = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) r += __low2float(acc[j]) + __high2float(acc[j]);
    out[i] = r;
}

// same fetch with the result returned as 2 x f16x2 (tex.3d.v2.f16x2.f32): no F2FP, half the return registers
__global__ void __launch_bounds__(128) fetch_tex_h2_kernel(Tex4 T, int steps, float* out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    __half2 acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = __float2half2_rn(0.f);
    for (int s = 0; s < steps; ++s) {
        float x, y, z; ray_pos(i, s, x, y, z);
#pragma unroll
        for (int g = 0; g < 4; ++g) {
            uint32_t a, b;
            asm("tex.3d.v2.f16x2.f32 {%0, %1}, [%2, {%3, %4, %5, %5}];" : "=r"(a), "=r"(b) : "l"(T.t[g]), "f"(x), "f"(y), "f"(z));
            acc[2 * g] = __hadd2(acc[2 * g], *reinterpret_cast<__half2*>(&a));
            acc[2 * g + 1] = __hadd2(acc[2 * g + 1], *reinterpret_cast<__half2*>(&b));
        }
    }
    float r = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) r += __low2float(acc[j]) + __high2float(acc[j]);
    out[i] = r;
}

__global__ void __launch_bounds__(128) fetch_ldg_kernel(const uint4* __restrict__ grid, int N, int steps, float* out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    __half2 tot[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) tot[j] = __float2half2_rn(0.f);
    for (int s = 0; s < steps; ++s) {
        float x, y, z; ray_pos(i, s, x, y, z);
        const float ux = x * N - 0.5f, uy = y * N - 0.5f, uz = z * N - 0.5f;
        const float flx = floorf(ux), fly = floorf(uy), flz = floorf(uz);
        const float fx = ux - flx, fy = uy - fly, fz = uz - flz;
        const int ix = int(flx), iy = int(fly), iz = int(flz);
        const int xs[2] = {min(max(ix, 0), N - 1), min(max(ix + 1, 0), N - 1)};
        const int ys[2] = {min(max(iy, 0), N - 1), min(max(iy + 1, 0), N - 1)};
        const int zs[2] = {min(max(iz, 0), N - 1), min(max(iz + 1, 0), N - 1)};
        const float wx[2] = {1.f - fx, fx}, wy[2] = {1.f - fy, fy}, wz[2] = {1.f - fz, fz};
        __half2 acc[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] = __float2half2_rn(0.f);
#pragma unroll
        for (int c = 0; c < 8; ++c) {
            const uint4* p = grid + (size_t)((zs[c >> 2] * N + ys[(c >> 1) & 1]) * N + xs[c & 1]) * 2;
            const uint4 a = __ldg(p), b = __ldg(p + 1);
            const __half2 w = __float2half2_rn(wz[c >> 2] * wy[(c >> 1) & 1] * wx[c & 1]);
            const uint32_t u[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[j] = __hfma2(w, *reinterpret_cast<const __half2*>(&u[j]), acc[j]);
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) tot[j] = __hadd2(tot[j], acc[j]);
    }
    float r = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) r += __low2float(tot[j]) + __high2float(tot[j]);
    out[i] = r;
}

// hybrid: channels 0-7 through the texture units, channels 8-15 through LDG.128 (are TMU filter rate and LSU return
// bandwidth independent limits?)
__global__ void __launch_bounds__(128) fetch_hybrid_kernel(Tex4 T, const uint4* __restrict__ grid, int N, int steps, float* out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    __half2 tot[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) tot[j] = __float2half2_rn(0.f);
    for (int s = 0; s < steps; ++s) {
        float x, y, z; ray_pos(i, s, x, y, z);
#pragma unroll
        for (int g = 0; g < 2; ++g) {
            const float4 v = tex3D<float4>(T.t[g], x, y, z);
            tot[2 * g] = __hadd2(tot[2 * g], __floats2half2_rn(v.x, v.y));
            tot[2 * g + 1] = __hadd2(tot[2 * g + 1], __floats2half2_rn(v.z, v.w));
        }
        const float ux = x * N - 0.5f, uy = y * N - 0.5f, uz = z * N - 0.5f;
        const float flx = floorf(ux), fly = floorf(uy), flz = floorf(uz);
        const float fx = ux - flx, fy = uy - fly, fz = uz - flz;
        const int ix = int(flx), iy = int(fly), iz = int(flz);
        const int xs[2] = {min(max(ix, 0), N - 1), min(max(ix + 1, 0), N - 1)};
        const int ys[2] = {min(max(iy, 0), N - 1), min(max(iy + 1, 0), N - 1)};
        const int zs[2] = {min(max(iz, 0), N - 1), min(max(iz + 1, 0), N - 1)};
        const float wx[2] = {1.f - fx, fx}, wy[2] = {1.f - fy, fy}, wz[2] = {1.f - fz, fz};
        __half2 acc[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[j] = __float2half2_rn(0.f);
#pragma unroll
        for (int c = 0; c < 8; ++c) {
            const uint4* p = grid + (size_t)((zs[c >> 2] * N + ys[(c >> 1) & 1]) * N + xs[c & 1]) * 2;
            const uint4 b = __ldg(p + 1);
            const __half2 w = __float2half2_rn(wz[c >> 2] * wy[(c >> 1) & 1] * wx[c & 1]);
            const uint32_t u[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[j] = __hfma2(w, *reinterpret_cast<const __half2*>(&u[j]), acc[j]);
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) tot[4 + j] = __hadd2(tot[4 + j], acc[j]);
    }
    float r = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) r += __low2float(tot[j]) + __high2float(tot[j]);
    out[i] = r;
}

// ---------------------------------------------------------------- 3. MUFU sin/cos throughput
__global__ void __launch_bounds__(128) mufu_kernel(int iters, float* out) {
    float a = threadIdx.x * 0.001f, s = 0.f;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int j = 0; j < 14; ++j) { s += __sinf(a + j) ; s += __cosf(a + j); }
        a += 0.01f;
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// 4. weight quantisation of the texture filter: a 2x2x2 half4 texture whose value is the x index -> result = alpha_q(a)
__global__ void texq_kernel(cudaTextureObject_t t, int n, float* out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float a = float(i) / float(n);            // wanted fractional weight
    const float x = (0.5f + a) / 2.f;               // normalised coordinate: xB = x*2 - 0.5 = a
    out[i] = tex3D<float4>(t, x, 0.25f, 0.25f).x;
}

int main() {
    cudaDeviceProp prop; CK(cudaGetDeviceProperties(&prop, 0));
    const int sms = prop.multiProcessorCount;
    int clk_khz = 0; cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0);
    printf("device %s, %d SMs, max clock %d MHz\n", prop.name, sms, clk_khz / 1000);
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    float ms;
    // 1
    unsigned long long* d_cyc; uint32_t* d_sink;
    CK(cudaMalloc(&d_cyc, sizeof(unsigned long long) * sms * 8)); CK(cudaMalloc(&d_sink, 4));
    for (int mode = 0; mode < 2; ++mode)
        for (int per_sm = 1; per_sm <= 8; per_sm *= 2) {
            const int iters = 20000, grid = sms * per_sm;
            for (int rep = 0; rep < 2; ++rep) {
                CK(cudaEventRecord(e0));
                if (mode == 0) tmem_kernel<0><<<grid, 128>>>(iters, d_cyc, d_sink); else tmem_kernel<1><<<grid, 128>>>(iters, d_cyc, d_sink);
                CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1)); CK(cudaGetLastError());
            }
            CK(cudaEventElapsedTime(&ms, e0, e1));
            unsigned long long c0; CK(cudaMemcpy(&c0, d_cyc, 8, cudaMemcpyDeviceToHost));
            const double bytes_per_cta_iter = 128.0 * 32 * 4;
            printf("tmem mode %d (%s) ctas/SM %d: %.1f cycles per x32 ld (CTA 0), %.1f B/clk/SM read, wall %.3f ms\n", mode,
                   mode ? "ld32+st16" : "ld32", per_sm, double(c0) / iters, bytes_per_cta_iter * per_sm / (double(c0) / iters), ms);
        }
    // 2
    const int N = 32;
    std::vector<__half> h(size_t(N) * N * N * 16);
    for (size_t i = 0; i < h.size(); ++i) h[i] = __float2half(float((i * 2654435761u) % 1000) / 1000.f);
    uint4* d_grid; CK(cudaMalloc(&d_grid, h.size() * 2)); CK(cudaMemcpy(d_grid, h.data(), h.size() * 2, cudaMemcpyHostToDevice));
    Tex4 T;
    for (int g = 0; g < 4; ++g) {
        cudaChannelFormatDesc fd = cudaCreateChannelDescHalf4();
        cudaArray_t arr; CK(cudaMalloc3DArray(&arr, &fd, make_cudaExtent(N, N, N)));
        std::vector<__half> sub(size_t(N) * N * N * 4);
        for (size_t v = 0; v < size_t(N) * N * N; ++v) for (int c = 0; c < 4; ++c) sub[v * 4 + c] = h[v * 16 + g * 4 + c];
        cudaMemcpy3DParms cp = {};
        cp.srcPtr = make_cudaPitchedPtr(sub.data(), N * 8, N, N); cp.dstArray = arr; cp.extent = make_cudaExtent(N, N, N); cp.kind = cudaMemcpyHostToDevice;
        CK(cudaMemcpy3D(&cp));
        cudaResourceDesc rd = {}; rd.resType = cudaResourceTypeArray; rd.res.array.array = arr;
        cudaTextureDesc td = {}; td.addressMode[0] = td.addressMode[1] = td.addressMode[2] = cudaAddressModeClamp;
        td.filterMode = cudaFilterModeLinear; td.readMode = cudaReadModeElementType; td.normalizedCoords = 1;
        CK(cudaCreateTextureObject(&T.t[g], &rd, &td, nullptr));
    }
    const int rays = 1 << 20, steps = 256;
    float* d_out; CK(cudaMalloc(&d_out, rays * 4));
    std::vector<float> o1(rays), o2(rays);
    for (int rep = 0; rep < 2; ++rep) {
        CK(cudaEventRecord(e0)); fetch_tex_kernel<<<rays / 128, 128>>>(T, steps, d_out); CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1)); CK(cudaGetLastError());
    }
    CK(cudaEventElapsedTime(&ms, e0, e1)); CK(cudaMemcpy(o1.data(), d_out, rays * 4, cudaMemcpyDeviceToHost));
    printf("fetch tex3D half4 x4: %.3f ms, %.3e samples/s, %.2f clk/sample/SM at max clock\n", ms, double(rays) * steps / (ms * 1e-3),
           ms * 1e-3 * clk_khz * 1e3 * sms / (double(rays) * steps));
    for (int rep = 0; rep < 2; ++rep) {
        CK(cudaEventRecord(e0)); fetch_tex_h2_kernel<<<rays / 128, 128>>>(T, steps, d_out); CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1)); CK(cudaGetLastError());
    }
    CK(cudaEventElapsedTime(&ms, e0, e1)); CK(cudaMemcpy(o2.data(), d_out, rays * 4, cudaMemcpyDeviceToHost));
    { double md2 = 0; for (int i = 0; i < rays; ++i) md2 = std::max(md2, double(std::fabs(o1[i] - o2[i])));
      printf("fetch tex3D half4 x4, f16x2 return: %.3f ms, %.3e samples/s, %.2f clk/sample/SM at max clock; max |diff| vs fp32 return + F2FP (sum of 256 steps x 16 ch) %.3e\n",
             ms, double(rays) * steps / (ms * 1e-3), ms * 1e-3 * clk_khz * 1e3 * sms / (double(rays) * steps), md2); }
    for (int rep = 0; rep < 2; ++rep) {
        CK(cudaEventRecord(e0)); fetch_ldg_kernel<<<rays / 128, 128>>>(d_grid, N, steps, d_out); CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1)); CK(cudaGetLastError());
    }
    CK(cudaEventElapsedTime(&ms, e0, e1)); CK(cudaMemcpy(o2.data(), d_out, rays * 4, cudaMemcpyDeviceToHost));
    printf("fetch LDG.128 x16 + HFMA2: %.3f ms, %.3e samples/s, %.2f clk/sample/SM at max clock\n", ms, double(rays) * steps / (ms * 1e-3),
           ms * 1e-3 * clk_khz * 1e3 * sms / (double(rays) * steps));
    for (int rep = 0; rep < 2; ++rep) {
        CK(cudaEventRecord(e0)); fetch_hybrid_kernel<<<rays / 128, 128>>>(T, d_grid, N, steps, d_out); CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1)); CK(cudaGetLastError());
    }
    CK(cudaEventElapsedTime(&ms, e0, e1));
    printf("fetch hybrid (2 tex3D + 8 LDG.128): %.3f ms, %.3e samples/s, %.2f clk/sample/SM at max clock\n", ms, double(rays) * steps / (ms * 1e-3),
           ms * 1e-3 * clk_khz * 1e3 * sms / (double(rays) * steps));
    double md = 0; for (int i = 0; i < rays; ++i) md = fmax(md, fabs(double(o1[i]) - o2[i]) / steps);
    printf("tex vs ldg: max |diff| per step (sum over 16 channels) %.3e\n", md);
    {   // 4
        cudaChannelFormatDesc fd = cudaCreateChannelDescHalf4();
        cudaArray_t arr; CK(cudaMalloc3DArray(&arr, &fd, make_cudaExtent(2, 2, 2)));
        std::vector<__half> sub(8 * 4);
        for (int v = 0; v < 8; ++v) for (int c = 0; c < 4; ++c) sub[v * 4 + c] = __float2half(float(v & 1));
        cudaMemcpy3DParms cp = {};
        cp.srcPtr = make_cudaPitchedPtr(sub.data(), 2 * 8, 2, 2); cp.dstArray = arr; cp.extent = make_cudaExtent(2, 2, 2); cp.kind = cudaMemcpyHostToDevice;
        CK(cudaMemcpy3D(&cp));
        cudaResourceDesc rd = {}; rd.resType = cudaResourceTypeArray; rd.res.array.array = arr;
        cudaTextureDesc td = {}; td.addressMode[0] = td.addressMode[1] = td.addressMode[2] = cudaAddressModeClamp;
        td.filterMode = cudaFilterModeLinear; td.readMode = cudaReadModeElementType; td.normalizedCoords = 1;
        cudaTextureObject_t t; CK(cudaCreateTextureObject(&t, &rd, &td, nullptr));
        const int n = 4096; float* d_q; CK(cudaMalloc(&d_q, n * 4));
        texq_kernel<<<n / 128, 128>>>(t, n, d_q); CK(cudaDeviceSynchronize());
        std::vector<float> q(n); CK(cudaMemcpy(q.data(), d_q, n * 4, cudaMemcpyDeviceToHost));
        // model candidates: round-to-nearest / truncation of a*256
        int bad_rn = 0, bad_tr = 0; double maxdev = 0;
        for (int i = 0; i < n; ++i) {
            const float a = float(i) / n;
            if (q[i] != floorf(a * 256.f + 0.5f) / 256.f) ++bad_rn;
            if (q[i] != floorf(a * 256.f) / 256.f) ++bad_tr;
            maxdev = fmax(maxdev, fabs(q[i] - a));
        }
        printf("texq: %d samples, mismatches vs round-to-nearest(a*256)/256: %d, vs trunc: %d, max |alpha_q - a| = %.5f\n", n, bad_rn, bad_tr, maxdev);
        printf("texq first values:"); for (int i = 0; i < 40; ++i) printf(" %.5f", q[i]); printf("\n");
    }
    // 3
    for (int rep = 0; rep < 2; ++rep) {
        CK(cudaEventRecord(e0)); mufu_kernel<<<sms * 16, 128>>>(4096, d_out); CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1)); CK(cudaGetLastError());
    }
    CK(cudaEventElapsedTime(&ms, e0, e1));
    printf("mufu 28 sin/cos per iter: %.3f ms, %.3e iters/s, %.2f clk/iter/SM at max clock\n", ms, double(sms) * 16 * 128 * 4096 / (ms * 1e-3),
           ms * 1e-3 * clk_khz * 1e3 / (16.0 * 128 * 4096));
    return 0;
}
