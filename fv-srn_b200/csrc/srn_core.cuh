// Fused SRN evaluation core for sm_100a: one warp evaluates 32 samples (one per lane) per call.
//
// Stands in for reference renderer/renderer_volume_tensorcores.cuh:736-1164
// (VolumeInterpolationTensorcores::eval) with a B200-first layout:
//   * samples are the M dimension of the MMA tiles (D[32 samples x H] = X[32 x K] * W^T), weights are the
//     small stationary B operand, kept in shared memory in *fragment order* (one conflict-free LDS.128 per
//     16x16 weight tile per lane), pinned for the CTA lifetime;
//   * layer-0 inputs (position, Fourier features, trilinear latent vector) are produced in registers by the
//     lane that owns the sample, staged once through a padded per-warp shared-memory tile and picked up with
//     ldmatrix; from then on activations never leave registers: the f16 accumulator fragment of
//     mma.m16n8k16 is bit-for-bit the A fragment of the next layer, so the whole MLP chains without
//     shared-memory round trips (the reference stores/reloads every layer, :922-936, 1018-1023);
//   * fp16 inputs and fp16 accumulation with the bias preloaded into the accumulator and K consumed in
//     chunks of 16 (Fourier block first, then latent chunks) - the same arithmetic contract as the
//     reference's wmma path (:856-910, 968-995);
//   * the latent grid is fp16, channel-last, fetched with 128-bit loads (8 corners x 32 B per 16 channels)
//     and blended with fp16x2 FMAs using fp32-computed corner weights.
#pragma once
#include <cuda_fp16.h>
#include <cstdint>

namespace fvsrn {

enum { D_ACT_RELU = 0, D_ACT_SINE = 1, D_ACT_SNAKE = 2, D_ACT_SNAKEALT = 3, D_ACT_NONE = 5 };
enum { D_OUT_DENSITY = 0, D_OUT_DENSITY_DIRECT = 1, D_OUT_RGBO = 2, D_OUT_RGBO_DIRECT = 3,
       // density + predicted gradient (4 last-layer outputs) / + curvature (6), renderer_volume_tensorcores.cuh:96-100
       D_OUT_DENSITY_GRADIENT = 4, D_OUT_DENSITY_GRADIENT_DIRECT = 5, D_OUT_DENSITY_GRADIENT_CUBIC = 6,
       D_OUT_DENSITY_CURVATURE = 7, D_OUT_DENSITY_CURVATURE_DIRECT = 8 };

struct GridGroup {       // 16 latent channels
    const __half* a;     // keyframe A, channel-last [Z][Y][X][Cgrid], already offset to this group's first channel
    const __half* b;     // keyframe B (== a when static)
    float frac;          // temporal lerp factor; 0 -> only A is fetched
    int voxel_stride;    // distance between voxels in uint4 (16 B) units
};

struct NetParams {
    const uint4* blob;   // fragment-ordered weights + biases + Fourier matrix, copied verbatim to smem
    int blob_bytes;
    int off_w0, off_wh, off_wlast, off_bias, off_fourier; // byte offsets inside the blob
    int num_hidden;      // H->H layers after layer 0
    int activation;
    float act_param;
    int output_mode;
    int has_time;        // PASS_TIME_TO_NETWORK
    float time_value;    // keyframe index passed as input channel 3 (reference :777-779)
    int dir_in_fourier;  // USE_DIRECTION == 2
    int gx, gy, gz;
    GridGroup groups[4];
    float box_min[3];
    float box_size[3];
    // tcgen05 path: latent grids as hardware-filtered fp16 RGBA 3-D textures (4 channels each; the reference's own
    // addressing: normalised coordinates, clamp, linear - volume_interpolation_network.cpp:470-504) and the Fourier
    // matrix as fp16x2 pairs in the kernel-parameter constant bank (index (k*NF + i)/2)
    unsigned long long tex_a[8], tex_b[8];
    uint32_t fourier2[192];
    int fourier_nerf;     // Fourier matrix has the NeRF layout (see SrnTc::build_row)
    uint32_t wlast2[64];  // density networks: last-layer weights as fp16x2 pairs (k, k+1) and its bias, for the in-thread dot product
    float blast;
    // byte-encoded latent grids (LatentGrid::Encoding, volume_interpolation_network.cpp:336-431): the grids stay 8-bit on the
    // device - unorm8 RGBA textures (tcgen05 / reference-exact kernels) or channel-last bytes (mma.sync kernels, groups[].a/b
    // then point to uint8) - and are decoded AFTER the trilinear filter as the reference does (EncodeGridValue,
    // renderer_volume_tensorcores.cuh:341-383): value = offset + scale * x (linear) or offset + scale * sqrt2 * erfinv((2 - 1e-4)(x - 0.5))
    int grid_encoding;    // 0 float (fp16 on the device), 1 byte-linear, 2 byte-gaussian
    float enc_offset_a[32], enc_scale_a[32];   // per latent channel, keyframe A
    float enc_offset_b[32], enc_scale_b[32];   // keyframe B (the reference decodes B with A's constants, :586-587; see make_net_params)
    // normals: GRADIENT_MODE of the reference (renderer_volume_tensorcores.cuh:132-134): 0 = none / predicted by
    // the network, 1 = central finite differences of the density with step fd_stepsize (world units, :1185-1197)
    int gradient_mode;
    float fd_stepsize;
    int off_wlast_rows;   // mma.sync blob: last layer once more as [cin][8 outputs] fp16 rows (reference-exact fp16 chains)
    // gradient_mode 2 (adjoint pass, kernels_exact.cuh): transposed weights as B fragments - hidden layers, the first H inputs of
    // layer 0, its 16-channel latent groups (behind the regular blob; blob_bytes includes them in this mode); fd_stepsize then
    // holds latentGridDifferencesStepSize
    int off_adj_wh, off_adj_w0f, off_adj_w0g;
};

// filtered byte-grid value -> latent value (EncodeGridValue). fast = reference-exact mode: the reference's fast-math operation order.
__device__ __forceinline__ float decode_grid_value(int encoding, float v, float offset, float scale) {
    if (encoding == 2) v = 1.4142135623730951f * erfinvf((2.0f - 1e-4f) * (v - 0.5f));
    return fmaf(v, scale, offset);
}

// ------------------------------------------------------------------ small PTX wrappers
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void ldmatrix_x4(uint32_t (&r)[4], uint32_t addr) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];\n"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
// D = A(16x16,row) * B(16x8,col) + C, all fp16, fp16 accumulate (HMMA.16816.F16)
__device__ __forceinline__ void mma16816(uint32_t (&d)[2], const uint32_t (&a)[4], uint32_t b0, uint32_t b1,
                                         uint32_t c0, uint32_t c1) {
    asm("mma.sync.aligned.m16n8k16.row.col.f16.f16.f16.f16 {%0,%1}, {%2,%3,%4,%5}, {%6,%7}, {%8,%9};\n"
        : "=r"(d[0]), "=r"(d[1])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1), "r"(c0), "r"(c1));
}
__device__ __forceinline__ uint4 lds128(uint32_t addr) {
    uint4 v; asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];\n" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
    return v;
}
__device__ __forceinline__ uint2 lds64(uint32_t addr) {
    uint2 v; asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];\n" : "=r"(v.x), "=r"(v.y) : "r"(addr));
    return v;
}
__device__ __forceinline__ unsigned short lds16(uint32_t addr) {
    unsigned short v; asm volatile("ld.shared.u16 %0, [%1];\n" : "=h"(v) : "r"(addr)); return v;
}
__device__ __forceinline__ uint32_t lds32(uint32_t addr) {
    uint32_t v; asm volatile("ld.shared.u32 %0, [%1];\n" : "=r"(v) : "r"(addr));
    return v;
}
__device__ __forceinline__ uint32_t h2u(__half2 h) { return *reinterpret_cast<uint32_t*>(&h); }
__device__ __forceinline__ __half2 u2h(uint32_t u) { return *reinterpret_cast<__half2*>(&u); }

// ------------------------------------------------------------------ hidden-layer activations (fp16, reference
// renderer/renderer_activations.cuh:39-341; same fp16 rounding points, sin/cos evaluated in fp32 by the SFU)
struct ActConst {
    __half2 p;      // Sine: param ; Snake: param
    __half2 q;      // Snake: 1/param ; SnakeAlt: 2*param
    float inv_q;    // SnakeAlt: 1 / float(half(2*param))
};
__device__ __forceinline__ ActConst make_act_const(int act, float param) {
    ActConst c;
    c.p = __float2half2_rn(param);
    c.q = (act == D_ACT_SNAKE) ? __float2half2_rn(1.0f / param) : __float2half2_rn(2.0f * param);
    c.inv_q = 1.0f / __low2float(c.q);
    return c;
}
__device__ __forceinline__ __half2 sin_h2(__half2 x) {
    const float2 f = __half22float2(x);
    return __floats2half2_rn(__sinf(f.x), __sinf(f.y));
}
__device__ __forceinline__ __half2 cos_h2(__half2 x) {
    const float2 f = __half22float2(x);
    return __floats2half2_rn(__cosf(f.x), __cosf(f.y));
}
template <int A, int B, int C, typename F>
__device__ __forceinline__ void for_each_reg(uint32_t (&r)[A][B][C], F f) {
#pragma unroll
    for (int a = 0; a < A; ++a)
#pragma unroll
        for (int b = 0; b < B; ++b)
#pragma unroll
            for (int c = 0; c < C; ++c) r[a][b][c] = f(r[a][b][c]);
}
template <int A, int B, int C>
__device__ __forceinline__ void activate(uint32_t (&r)[A][B][C], int act, const ActConst& c) {
    if (act == D_ACT_RELU) {
        const __half2 z = __float2half2_rn(0.f);
        for_each_reg(r, [&](uint32_t v) { return h2u(__hmax2(u2h(v), z)); });
    } else if (act == D_ACT_SINE) {          // hsin(v * p)
        for_each_reg(r, [&](uint32_t v) { return h2u(sin_h2(__hmul2(u2h(v), c.p))); });
    } else if (act == D_ACT_SNAKE) {         // v + (1/p) * sin(p v)^2
        for_each_reg(r, [&](uint32_t u) {
            const __half2 v = u2h(u);
            const __half2 s = sin_h2(__hmul2(c.p, v));
            return h2u(__hadd2(v, __hmul2(c.q, __hmul2(s, s))));
        });
    } else if (act == D_ACT_SNAKEALT) {      // (v + 1 - cos(2p v)) / (2p)
        const __half2 one = __float2half2_rn(1.f);
        for_each_reg(r, [&](uint32_t u) {
            const __half2 v = u2h(u);
            const __half2 x0 = cos_h2(__hmul2(c.q, v));
            const float2 x1 = __half22float2(__hsub2(__hadd2(v, one), x0));
            return h2u(__floats2half2_rn(x1.x * c.inv_q, x1.y * c.inv_q)); // __hdiv = fp32 mul by rcp, then round
        });
    }
}

// ------------------------------------------------------------------ the network
template <int H16, int GC16, bool DIR>
struct Srn {
    static constexpr int H = 16 * H16;            // hidden channels
    static constexpr int GC = 16 * GC16;          // latent channels
    static constexpr int K0 = H + GC;             // inputs of layer 0
    static constexpr int FOFF = DIR ? 8 : 4;      // first Fourier channel
    static constexpr int NF = (H - FOFF) / 2;     // Fourier features (2*NF + FOFF == H, reference :142-152)
    static constexpr int KT0 = K0 / 16;
    static constexpr int KTH = H / 16;
    static constexpr int NT = H / 8;              // n-tiles of 8
    static constexpr int NT2 = H / 16;            // n-tile pairs
    static constexpr int ROW_CHUNKS = K0 / 8;     // 16-byte chunks per staged row
    // odd number of chunks per row => the 8 row addresses of an ldmatrix phase and the 8 lanes of an
    // STS.128 phase fall into 8 distinct 16-byte bank groups (conflict-free)
    static constexpr int RS_CHUNKS = (ROW_CHUNKS % 2 == 0) ? ROW_CHUNKS + 1 : ROW_CHUNKS;
    static constexpr int STAGE_BYTES = 32 * RS_CHUNKS * 16;
    static constexpr int ACC = 2 * NT * 2;        // accumulator registers per lane: [mt][nt][2]

    static __host__ __device__ constexpr int w0_bytes() { return K0 * H * 2; }
    static __host__ __device__ constexpr int wh_bytes() { return H * H * 2; }
    static __host__ __device__ constexpr int wlast_bytes() { return KTH * 32 * 8; }

    // one dense layer on the tensor cores: acc_out = bias + A * W^T, A given as fragments
    // a(mt,kt) is supplied by a functor so that layer 0 (ldmatrix) and hidden layers (registers) share the code
    template <int KT, typename AFrag>
    static __device__ __forceinline__ void dense(uint32_t (&out)[2][NT][2], uint32_t w_addr, uint32_t bias_addr,
                                                 int lane, AFrag afrag) {
        uint32_t bias[NT];
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) bias[nt] = lds32(bias_addr + (nt * 8 + (lane & 3) * 2) * 2);
#pragma unroll
        for (int kt = 0; kt < KT; ++kt) {
            uint32_t a[2][4];
            afrag(kt, a);
#pragma unroll
            for (int ntp = 0; ntp < NT2; ++ntp) {
                const uint4 b = lds128(w_addr + ((kt * NT2 + ntp) * 32 + lane) * 16);
#pragma unroll
                for (int mt = 0; mt < 2; ++mt) {
                    if (kt == 0) {
                        mma16816(out[mt][2 * ntp], a[mt], b.x, b.y, bias[2 * ntp], bias[2 * ntp]);
                        mma16816(out[mt][2 * ntp + 1], a[mt], b.z, b.w, bias[2 * ntp + 1], bias[2 * ntp + 1]);
                    } else {
                        mma16816(out[mt][2 * ntp], a[mt], b.x, b.y, out[mt][2 * ntp][0], out[mt][2 * ntp][1]);
                        mma16816(out[mt][2 * ntp + 1], a[mt], b.z, b.w, out[mt][2 * ntp + 1][0], out[mt][2 * ntp + 1][1]);
                    }
                }
            }
        }
    }

    // Layer-0 inputs of the sample owned by this thread, packed fp16x2:
    // [px,py,pz,time|0,(dx,dy,dz,0),cos x NF,sin x NF,latent x GC]   (reference :768-808, 839-910)
    // ch3_one: put 1.0 into channel 3 when the network has no time input (the tcgen05 path folds the bias there).
    static __device__ __forceinline__ void build_row(const NetParams& P, uint32_t sblob, float px, float py, float pz,
                                                     float dx, float dy, float dz, bool ch3_one, uint32_t (&row)[K0 / 2]) {
        const __half hx = __float2half_rn(px), hy = __float2half_rn(py), hz = __float2half_rn(pz);
        row[0] = h2u(__halves2half2(hx, hy));
        row[1] = h2u(__halves2half2(hz, P.has_time ? __float2half_rn(P.time_value) : __float2half_rn(ch3_one ? 1.f : 0.f)));
        __half2 d2x, d2y, d2z;
        if (DIR) {
            const __half hdx = __float2half_rn(dx), hdy = __float2half_rn(dy), hdz = __float2half_rn(dz);
            row[2] = h2u(__halves2half2(hdx, hdy));
            row[3] = h2u(__halves2half2(hdz, __float2half_rn(0.f)));
            d2x = __half2half2(hdx); d2y = __half2half2(hdy); d2z = __half2half2(hdz);
        }
        {   // Fourier features, reference :795-807: phase accumulated in fp16 (hmul, hfma, hfma), then cos / sin
            const __half2 x2 = __half2half2(hx), y2 = __half2half2(hy), z2 = __half2half2(hz);
            const uint32_t f_addr = sblob + P.off_fourier;
#pragma unroll
            for (int i = 0; i < NF; i += 2) {
                __half2 c = __hmul2(x2, u2h(lds32(f_addr + (i + 0 * NF) * 2)));
                c = __hfma2(y2, u2h(lds32(f_addr + (i + 1 * NF) * 2)), c);
                c = __hfma2(z2, u2h(lds32(f_addr + (i + 2 * NF) * 2)), c);
                if (DIR) {
                    if (P.dir_in_fourier) {
                        c = __hfma2(d2x, u2h(lds32(f_addr + (i + 3 * NF) * 2)), c);
                        c = __hfma2(d2y, u2h(lds32(f_addr + (i + 4 * NF) * 2)), c);
                        c = __hfma2(d2z, u2h(lds32(f_addr + (i + 5 * NF) * 2)), c);
                    }
                }
                const float2 cf = __half22float2(c);
                float s0, c0, s1, c1;
                __sincosf(cf.x, &s0, &c0);
                __sincosf(cf.y, &s1, &c1);
                row[FOFF / 2 + i / 2] = h2u(__floats2half2_rn(c0, c1));
                row[FOFF / 2 + NF / 2 + i / 2] = h2u(__floats2half2_rn(s0, s1));
            }
        }
        if (GC16 > 0) {
            // trilinear latent fetch. Sampling position p*N - 0.5 with border clamp == tex3D(normalised, clamp, linear)
            // == grid_sample(align_corners=False, padding_mode='border') (reference :573-605, volume_interpolation_network.cpp:470-504)
            const float ux = __fmaf_rn(px, float(P.gx), -0.5f), uy = __fmaf_rn(py, float(P.gy), -0.5f), uz = __fmaf_rn(pz, float(P.gz), -0.5f);
            const float flx = floorf(ux), fly = floorf(uy), flz = floorf(uz);
            const float fx = ux - flx, fy = uy - fly, fz = uz - flz;
            const int ix = int(flx), iy = int(fly), iz = int(flz);
            const int xs[2] = {min(max(ix, 0), P.gx - 1), min(max(ix + 1, 0), P.gx - 1)};
            const int ys[2] = {min(max(iy, 0), P.gy - 1), min(max(iy + 1, 0), P.gy - 1)};
            const int zs[2] = {min(max(iz, 0), P.gz - 1), min(max(iz + 1, 0), P.gz - 1)};
            const float wx[2] = {1.f - fx, fx}, wy[2] = {1.f - fy, fy}, wz[2] = {1.f - fz, fz};
#pragma unroll
            for (int g = 0; g < GC16; ++g) {
                const GridGroup G = P.groups[g];
                if (P.grid_encoding != 0) {     // 8-bit grid: blend the normalised texels in fp32, decode afterwards, then the temporal lerp
                    float va[16], vb[16];
#pragma unroll
                    for (int j = 0; j < 16; ++j) va[j] = vb[j] = 0.f;
                    const int nsrc8 = (G.frac > 0.f) ? 2 : 1;
#pragma unroll 1
                    for (int s8 = 0; s8 < nsrc8; ++s8) {
                        const uint4* base8 = reinterpret_cast<const uint4*>(s8 == 0 ? G.a : G.b);
                        float (&dst)[16] = s8 == 0 ? va : vb;
#pragma unroll
                        for (int c = 0; c < 8; ++c) {
                            const int zi = zs[c >> 2], yi = ys[(c >> 1) & 1], xi = xs[c & 1];
                            const uint4 t = __ldg(base8 + (size_t)((zi * P.gy + yi) * P.gx + xi) * G.voxel_stride);
                            const float wc = __fmul_rn(__fmul_rn(wz[c >> 2], wy[(c >> 1) & 1]), wx[c & 1]);
                            const uint32_t u[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
                            for (int j = 0; j < 16; ++j)
                                dst[j] = fmaf(wc, float((u[j >> 2] >> (8 * (j & 3))) & 0xffu) * (1.0f / 255.0f), dst[j]);
                        }
                    }
#pragma unroll
                    for (int j = 0; j < 16; ++j) {
                        float v = decode_grid_value(P.grid_encoding, va[j], P.enc_offset_a[16 * g + j], P.enc_scale_a[16 * g + j]);
                        if (nsrc8 == 2) {
                            const float w = decode_grid_value(P.grid_encoding, vb[j], P.enc_offset_b[16 * g + j], P.enc_scale_b[16 * g + j]);
                            v = fmaf(G.frac, w - v, v);
                        }
                        va[j] = v;
                    }
#pragma unroll
                    for (int j = 0; j < 8; ++j) row[H / 2 + g * 8 + j] = h2u(__floats2half2_rn(va[2 * j], va[2 * j + 1]));
                    continue;
                }
                __half2 acc[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[j] = __float2half2_rn(0.f);
                const int nsrc = (G.frac > 0.f) ? 2 : 1; // warp-uniform
#pragma unroll 1
                for (int s = 0; s < nsrc; ++s) {
                    const uint4* base = reinterpret_cast<const uint4*>(s == 0 ? G.a : G.b);
                    const float ws = (nsrc == 1) ? 1.f : (s == 0 ? 1.f - G.frac : G.frac);
                    uint4 v[8][2];
                    __half2 w2[8];
#pragma unroll
                    for (int c = 0; c < 8; ++c) {
                        const int zi = zs[c >> 2], yi = ys[(c >> 1) & 1], xi = xs[c & 1];
                        const uint4* ptr = base + (size_t)((zi * P.gy + yi) * P.gx + xi) * G.voxel_stride;
                        v[c][0] = __ldg(ptr);
                        v[c][1] = __ldg(ptr + 1);
                        w2[c] = __float2half2_rn(__fmul_rn(__fmul_rn(__fmul_rn(wz[c >> 2], wy[(c >> 1) & 1]), wx[c & 1]), ws));
                    }
#pragma unroll
                    for (int c = 0; c < 8; ++c) {
                        const uint32_t u[8] = {v[c][0].x, v[c][0].y, v[c][0].z, v[c][0].w, v[c][1].x, v[c][1].y, v[c][1].z, v[c][1].w};
#pragma unroll
                        for (int j = 0; j < 8; ++j)
                            acc[j] = __hfma2(w2[c], u2h(u[j]), acc[j]);
                    }
                }
#pragma unroll
                for (int j = 0; j < 8; ++j) row[H / 2 + g * 8 + j] = h2u(acc[j]);
            }
        }
    }

    // ------------------------------------------------------------------------------------------------------------------
    // Adjoint pass (the reference's GRADIENT_MODE 2, renderer_volume_tensorcores.cuh:1205-1530) in this family's arithmetic:
    // d density / d position of the sample owned by each lane from one backward pass - transposed-weight MMAs (fp16
    // accumulators from zero), activation adjoints in packed fp16, the latent grid differentiated by central differences with
    // step P.fd_stepsize in normalised grid coordinates (LoadVolumetricFeaturesAdjoint, :609-735), the Fourier adjoint in fp32.
    // The reference-arithmetic version (bit-identical to the reference kernel) is SrnExact::adjoint in kernels_exact.cuh.
    static __host__ __device__ constexpr int pre_words(int num_hidden) { return (1 + num_hidden) * ACC * 32; }
    static __device__ __forceinline__ void keep_frag(uint32_t* pre, int layer, const uint32_t (&f)[2][NT][2], int lane) {
#pragma unroll
        for (int mt = 0; mt < 2; ++mt)
#pragma unroll
            for (int nt = 0; nt < NT; ++nt)
#pragma unroll
                for (int j = 0; j < 2; ++j) pre[(layer * ACC + (mt * NT + nt) * 2 + j) * 32 + lane] = f[mt][nt][j];
    }
    static __device__ __forceinline__ void act_adjoint(uint32_t (&z)[2][NT][2], const uint32_t* pre, int layer, int lane, int act, const ActConst& c) {
        const __half2 zero = __float2half2_rn(0.f), one = __float2half2_rn(1.f);
        const __half2 two_p = __hadd2(c.p, c.p), inv_q = __float2half2_rn(c.inv_q);
#pragma unroll
        for (int mt = 0; mt < 2; ++mt)
#pragma unroll
            for (int nt = 0; nt < NT; ++nt)
#pragma unroll
                for (int j = 0; j < 2; ++j) {
                    const __half2 x = u2h(pre[(layer * ACC + (mt * NT + nt) * 2 + j) * 32 + lane]);
                    __half2 v = u2h(z[mt][nt][j]);
                    if (act == D_ACT_RELU) v = __hmul2(v, __hgt2(x, zero));                                   // x > 0 ? z : 0
                    else if (act == D_ACT_SINE) v = __hmul2(v, __hmul2(c.p, cos_h2(__hmul2(x, c.p))));         // z p cos(p x)
                    else if (act == D_ACT_SNAKE) v = __hmul2(v, __hadd2(one, sin_h2(__hmul2(two_p, x))));      // z (1 + sin(2p x))
                    else if (act == D_ACT_SNAKEALT) v = __hmul2(v, __hadd2(sin_h2(__hmul2(c.q, x)), inv_q));   // z (sin(2p x) + 1/(2p))
                    z[mt][nt][j] = h2u(v);
                }
    }
    // out(32 x 8*NTL) = X(32 x H) * B, B = transposed weights as fragments (api.cu build_blob, adjoint sections)
    template <int NTL>
    static __device__ __forceinline__ void dense_t(uint32_t (&out)[2][NTL][2], uint32_t w_addr, int lane, const uint32_t (&x)[2][NT][2]) {
#pragma unroll
        for (int kt = 0; kt < KTH; ++kt) {
            uint32_t a[2][4];
#pragma unroll
            for (int mt = 0; mt < 2; ++mt) {
                a[mt][0] = x[mt][2 * kt][0]; a[mt][1] = x[mt][2 * kt][1];
                a[mt][2] = x[mt][2 * kt + 1][0]; a[mt][3] = x[mt][2 * kt + 1][1];
            }
#pragma unroll
            for (int ntp = 0; ntp < NTL / 2; ++ntp) {
                const uint4 b = lds128(w_addr + ((kt * (NTL / 2) + ntp) * 32 + lane) * 16);
#pragma unroll
                for (int mt = 0; mt < 2; ++mt) {
                    if (kt == 0) {
                        mma16816(out[mt][2 * ntp], a[mt], b.x, b.y, 0u, 0u);
                        mma16816(out[mt][2 * ntp + 1], a[mt], b.z, b.w, 0u, 0u);
                    } else {
                        mma16816(out[mt][2 * ntp], a[mt], b.x, b.y, out[mt][2 * ntp][0], out[mt][2 * ntp][1]);
                        mma16816(out[mt][2 * ntp + 1], a[mt], b.z, b.w, out[mt][2 * ntp + 1][0], out[mt][2 * ntp + 1][1]);
                    }
                }
            }
        }
    }
    template <int NTL>
    static __device__ __forceinline__ void frag_to_rows(uint4* stage, const uint32_t (&f)[2][NTL][2], int lane) {
        uint32_t* st32 = reinterpret_cast<uint32_t*>(stage);
#pragma unroll
        for (int mt = 0; mt < 2; ++mt)
#pragma unroll
            for (int nt = 0; nt < NTL; ++nt)
#pragma unroll
                for (int j = 0; j < 2; ++j)
                    st32[(mt * 16 + (lane >> 2) + 8 * j) * (RS_CHUNKS * 4) + nt * 4 + (lane & 3)] = f[mt][nt][j];
    }
    // the 16 latent channels of group g at one position (float grids; fp16 weights and FMAs as in build_row), as floats
    static __device__ __forceinline__ void latent16(const NetParams& P, int g, float px, float py, float pz, float (&v)[16]) {
        const float ux = __fmaf_rn(px, float(P.gx), -0.5f), uy = __fmaf_rn(py, float(P.gy), -0.5f), uz = __fmaf_rn(pz, float(P.gz), -0.5f);
        const float flx = floorf(ux), fly = floorf(uy), flz = floorf(uz);
        const float fx = ux - flx, fy = uy - fly, fz = uz - flz;
        const int ix = int(flx), iy = int(fly), iz = int(flz);
        const int xs[2] = {min(max(ix, 0), P.gx - 1), min(max(ix + 1, 0), P.gx - 1)};
        const int ys[2] = {min(max(iy, 0), P.gy - 1), min(max(iy + 1, 0), P.gy - 1)};
        const int zs[2] = {min(max(iz, 0), P.gz - 1), min(max(iz + 1, 0), P.gz - 1)};
        const float wx[2] = {1.f - fx, fx}, wy[2] = {1.f - fy, fy}, wz[2] = {1.f - fz, fz};
        const GridGroup G = P.groups[g];
        __half2 acc[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] = __float2half2_rn(0.f);
        const int nsrc = (G.frac > 0.f) ? 2 : 1;
#pragma unroll 1
        for (int s = 0; s < nsrc; ++s) {
            const uint4* base = reinterpret_cast<const uint4*>(s == 0 ? G.a : G.b);
            const float ws = (nsrc == 1) ? 1.f : (s == 0 ? 1.f - G.frac : G.frac);
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                const uint4* ptr = base + (size_t)((zs[c >> 2] * P.gy + ys[(c >> 1) & 1]) * P.gx + xs[c & 1]) * G.voxel_stride;
                const uint4 v0 = __ldg(ptr), v1 = __ldg(ptr + 1);
                const __half2 w2 = __float2half2_rn(__fmul_rn(__fmul_rn(__fmul_rn(wz[c >> 2], wy[(c >> 1) & 1]), wx[c & 1]), ws));
                const uint32_t u[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[j] = __hfma2(w2, u2h(u[j]), acc[j]);
            }
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) { const float2 f = __half22float2(acc[j]); v[2 * j] = f.x; v[2 * j + 1] = f.y; }
    }
    // all 32 lanes call, after forward<true> of the same positions. raw0 = the raw density output of this lane's sample.
    static __device__ __forceinline__ float3 adjoint(const NetParams& P, uint32_t sblob, uint4* stage, const uint32_t* pre, const ActConst& ac,
                                                     float px, float py, float pz, float raw0) {
        const int lane = threadIdx.x & 31;
        // adjoint of the output parametrisation (:1233-1246): sigmoid'(raw) for the sigmoid modes, else 1
        float adj_out = 1.f;
        if (P.output_mode == D_OUT_DENSITY || P.output_mode == D_OUT_DENSITY_GRADIENT) {
            const float ev = __expf(raw0);
            adj_out = __fdividef(ev, (ev + 1.f) * (ev + 1.f));
        }
        const __half adj_h = __float2half_rn(adj_out);
        const uint32_t adj_bits = uint32_t(__half_as_ushort(adj_h));
        // adjoint of the last layer (:1248-1262), built directly as accumulator fragments: sample mt*16 + lane/4 + 8j, channel pair nt*4 + lane%4
        uint32_t z[2][NT][2];
#pragma unroll
        for (int mt = 0; mt < 2; ++mt)
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                const __half a = __ushort_as_half((unsigned short)__shfl_sync(0xffffffffu, adj_bits, mt * 16 + (lane >> 2) + 8 * j));
                const __half2 a2 = __halves2half2(a, a);
#pragma unroll
                for (int nt = 0; nt < NT; ++nt) z[mt][nt][j] = h2u(__hmul2(u2h(P.wlast2[nt * 4 + (lane & 3)]), a2));
            }
#pragma unroll 1
        for (int l = P.num_hidden - 1; l >= 0; --l) {
            act_adjoint(z, pre, 1 + l, lane, P.activation, ac);
            uint32_t nxt[2][NT][2];
            dense_t<NT>(nxt, sblob + P.off_adj_wh + l * wh_bytes(), lane, z);
#pragma unroll
            for (int mt = 0; mt < 2; ++mt)
#pragma unroll
                for (int nt = 0; nt < NT; ++nt) { z[mt][nt][0] = nxt[mt][nt][0]; z[mt][nt][1] = nxt[mt][nt][1]; }
        }
        act_adjoint(z, pre, 0, lane, P.activation, ac);
        float3 out = make_float3(0.f, 0.f, 0.f);
        const float h = P.fd_stepsize;
        const float s = __fdividef(1.f, 2.f * h);
#pragma unroll
        for (int g = 0; g < GC16; ++g) {
            uint32_t c16[2][2][2];
            dense_t<2>(c16, sblob + P.off_adj_w0g + g * (H * 16 * 2), lane, z);
            __syncwarp();
            frag_to_rows<2>(stage, c16, lane);
            __syncwarp();
            const uint4 r0 = stage[lane * RS_CHUNKS], r1 = stage[lane * RS_CHUNKS + 1];
            const uint32_t u[8] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w};
            float adj[16];
#pragma unroll
            for (int k = 0; k < 8; ++k) { const float2 f = __half22float2(u2h(u[k])); adj[2 * k] = f.x; adj[2 * k + 1] = f.y; }
#pragma unroll 1
            for (int axis = 0; axis < 3; ++axis) {
                float hi[16], lo[16];
                latent16(P, g, px + (axis == 0 ? h : 0.f), py + (axis == 1 ? h : 0.f), pz + (axis == 2 ? h : 0.f), hi);
                latent16(P, g, px - (axis == 0 ? h : 0.f), py - (axis == 1 ? h : 0.f), pz - (axis == 2 ? h : 0.f), lo);
                float d = 0.f;
#pragma unroll
                for (int k = 0; k < 16; ++k) d = fmaf(adj[k], s * (hi[k] - lo[k]), d);
                if (axis == 0) out.x += d; else if (axis == 1) out.y += d; else out.z += d;
            }
        }
        {
            uint32_t zf[2][NT][2];
            dense_t<NT>(zf, sblob + P.off_adj_w0f, lane, z);
            __syncwarp();
            frag_to_rows<NT>(stage, zf, lane);
            __syncwarp();
        }
        const __half* row = reinterpret_cast<const __half*>(stage + lane * RS_CHUNKS);
        out.x += __half2float(row[0]); out.y += __half2float(row[1]); out.z += __half2float(row[2]);
        const __half hx = __float2half_rn(px), hy = __float2half_rn(py), hz = __float2half_rn(pz);
        const uint32_t f_addr = sblob + P.off_fourier;
#pragma unroll 2
        for (int i = 0; i < NF; ++i) {
            const __half f0 = __ushort_as_half(lds16(f_addr + (i + 0 * NF) * 2)), f1 = __ushort_as_half(lds16(f_addr + (i + 1 * NF) * 2)),
                         f2 = __ushort_as_half(lds16(f_addr + (i + 2 * NF) * 2));
            __half c = __hmul(hx, f0);
            c = __hfma(hy, f1, c);
            c = __hfma(hz, f2, c);
            float sn, cs;
            __sincosf(__half2float(c), &sn, &cs);
            const float adj_c = __half2float(row[FOFF + NF + i]) * cs - __half2float(row[FOFF + i]) * sn;
            out.x = fmaf(__half2float(f0), adj_c, out.x); out.y = fmaf(__half2float(f1), adj_c, out.y); out.z = fmaf(__half2float(f2), adj_c, out.z);
        }
        __syncwarp();
        return out;
    }

    // Evaluate the network for the sample owned by this lane. All 32 lanes must call (warp-synchronous).
    // (px,py,pz) is the position in the unit cube, (dx,dy,dz) the ray direction (DIR only).
    // out[0..3]: raw last-layer outputs (before the output parametrisation).
    // KEEP: the pre-activation accumulators of every activated layer go to `pre` (fragment order, [(layer * ACC + reg) * 32 + lane])
    // for the adjoint pass below
    template <bool KEEP = false>
    static __device__ __forceinline__ void forward(const NetParams& P, uint32_t sblob, uint4* stage, const ActConst& ac,
                                                   float px, float py, float pz, float dx, float dy, float dz,
                                                   float (&out)[4], uint32_t* pre = nullptr) {
        const int lane = threadIdx.x & 31;
        __syncwarp();
        uint32_t row[K0 / 2];
        build_row(P, sblob, px, py, pz, dx, dy, dz, false, row);
        // ---------------- stage the 32 x K0 tile (row = lane) and pick it up as A fragments
#pragma unroll
        for (int j = 0; j < ROW_CHUNKS; ++j)
            stage[lane * RS_CHUNKS + j] = make_uint4(row[4 * j], row[4 * j + 1], row[4 * j + 2], row[4 * j + 3]);
        __syncwarp();
        const uint32_t stage_addr = smem_u32(stage);
        const uint32_t bias_addr = sblob + P.off_bias;
        uint32_t acc[2][NT][2];
        {
            const int lrow = (lane & 7) + ((lane >> 3) & 1) * 8;
            const int lcol = lane >> 4;
            dense<KT0>(acc, sblob + P.off_w0, bias_addr, lane, [&](int kt, uint32_t (&a)[2][4]) {
#pragma unroll
                for (int mt = 0; mt < 2; ++mt)
                    ldmatrix_x4(a[mt], stage_addr + ((mt * 16 + lrow) * RS_CHUNKS + kt * 2 + lcol) * 16);
            });
        }
        if constexpr (KEEP) keep_frag(pre, 0, acc, lane);
        activate(acc, P.activation, ac);
        // ---------------- hidden layers: accumulator fragments are the next A fragments
#pragma unroll 1
        for (int l = 0; l < P.num_hidden; ++l) {
            uint32_t nxt[2][NT][2];
            dense<KTH>(nxt, sblob + P.off_wh + l * wh_bytes(), bias_addr + (1 + l) * H * 2, lane,
                       [&](int kt, uint32_t (&a)[2][4]) {
#pragma unroll
                           for (int mt = 0; mt < 2; ++mt) {
                               a[mt][0] = acc[mt][2 * kt][0]; a[mt][1] = acc[mt][2 * kt][1];
                               a[mt][2] = acc[mt][2 * kt + 1][0]; a[mt][3] = acc[mt][2 * kt + 1][1];
                           }
                       });
            if constexpr (KEEP) keep_frag(pre, 1 + l, nxt, lane);
            activate(nxt, P.activation, ac);
#pragma unroll
            for (int mt = 0; mt < 2; ++mt)
#pragma unroll
                for (int nt = 0; nt < NT; ++nt) { acc[mt][nt][0] = nxt[mt][nt][0]; acc[mt][nt][1] = nxt[mt][nt][1]; }
        }
        // ---------------- last layer (H -> up to 8 outputs, padded with zero weights), also on the tensor cores
        uint32_t o[2][2];
        {
            const uint32_t wl = sblob + P.off_wlast;
            const uint32_t bl = lds32(bias_addr + ((1 + P.num_hidden) * H + (lane & 3) * 2) * 2);
#pragma unroll
            for (int kt = 0; kt < KTH; ++kt) {
                const uint2 b = lds64(wl + (kt * 32 + lane) * 8);
#pragma unroll
                for (int mt = 0; mt < 2; ++mt) {
                    const uint32_t a[4] = {acc[mt][2 * kt][0], acc[mt][2 * kt][1], acc[mt][2 * kt + 1][0], acc[mt][2 * kt + 1][1]};
                    if (kt == 0) mma16816(o[mt], a, b.x, b.y, bl, bl);
                    else mma16816(o[mt], a, b.x, b.y, o[mt][0], o[mt][1]);
                }
            }
        }
        // sample s = lane lives in m-tile s/16, row r = s%16: fragment register r/8 of lane 4*(r%8) + col/2
        {
            const int r = lane & 15;
            const int src = (r & 7) * 4;
            const bool hi = (r >> 3) != 0, m1 = (lane >> 4) != 0;
            const uint32_t v00 = __shfl_sync(0xffffffffu, o[0][0], src), v01 = __shfl_sync(0xffffffffu, o[0][1], src);
            const uint32_t v10 = __shfl_sync(0xffffffffu, o[1][0], src), v11 = __shfl_sync(0xffffffffu, o[1][1], src);
            const float2 f = __half22float2(u2h(m1 ? (hi ? v11 : v10) : (hi ? v01 : v00)));
            out[0] = f.x; out[1] = f.y; out[2] = 0.f; out[3] = 0.f;
            if (P.output_mode >= D_OUT_RGBO) {
                const uint32_t w00 = __shfl_sync(0xffffffffu, o[0][0], src + 1), w01 = __shfl_sync(0xffffffffu, o[0][1], src + 1);
                const uint32_t w10 = __shfl_sync(0xffffffffu, o[1][0], src + 1), w11 = __shfl_sync(0xffffffffu, o[1][1], src + 1);
                const float2 g = __half22float2(u2h(m1 ? (hi ? w11 : w10) : (hi ? w01 : w00)));
                out[2] = g.x; out[3] = g.y;
            }
        }
    }
};

// output parametrisation, reference :1054-1072, 1149-1158 and renderer_activations.cuh (Sigmoid / Softplus, fp32)
__device__ __forceinline__ float sigmoidf_(float v) { return 1.0f / (1.0f + __expf(-v)); }
__device__ __forceinline__ float softplusf_(float v) { return v > 20.f ? v : __logf(1.0f + __expf(v)); }

} // namespace fvsrn
