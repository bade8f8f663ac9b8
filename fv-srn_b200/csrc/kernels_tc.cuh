// tcgen05 / TMEM versions of the two kernels of the hot path (fast mode, see srn_tc.cuh):
//   dvr_tc_kernel   <- ImageEvaluatorSimpleKernel + RayEvaluationSteppingDvr::eval + TF + Blending
//   eval_tc_kernel  <- EvaluateNoBatches
// One CTA = 4 compute warps (128 threads = 128 rays, a 16x8 pixel tile, each warp an 8x4 sub-tile) + 1 MMA warp.
// The rays are marched in lock-step: one SRN sample per compute thread per step = one M=128 UMMA tile per step, two
// steps ("slots") in flight. Compute warps never wait for each other: they publish their operand rows with an
// mbarrier arrive and move on to the other slot; the elected thread of the MMA warp waits for the four warps of a
// slot, issues that layer's tcgen05.mma chain and commits to the slot's completion mbarrier. Persistent CTAs pull
// tiles from an atomic queue; several CTAs are resident per SM.
#pragma once
#include "kernels.cuh"
#include "srn_tc.cuh"

namespace fvsrn {

// SPEC (template parameter of the DVR kernel and of SrnTc): compile-time specialisation of launch-uniform parameters.
// The march loop tests them per sample or per trip; as run-time values they cost branches and keep the compiler from
// scheduling across them. Measured on C3 (A/B on one B200, r02c): 5.49 ms generic, 4.69 ms with them frozen.
//   0  generic: everything at run time
//   1  the common DVR configuration: density:direct network without time input, piecewise TF shared by all views,
//      Beer-Lambert blending, early termination on, one sample per pixel, static latent grid (no temporal interpolation)
//   2  as 1 with temporally interpolated latent grids (both keyframes always fetched and blended)
template <int SPEC>
struct TcSpec {
    static constexpr bool FAST = SPEC != 0;
    static __device__ __forceinline__ int output_mode(const NetParams& P) { return FAST ? int(D_OUT_DENSITY_DIRECT) : P.output_mode; }
    static __device__ __forceinline__ int has_time(const NetParams& P) { return FAST ? 0 : P.has_time; }
    static __device__ __forceinline__ int tf_kind(const DvrParams& D) { return FAST ? 0 : D.tf_kind; }
    static __device__ __forceinline__ int tf_batch(const DvrParams& D) { return FAST ? 1 : D.tf_batch; }
    static __device__ __forceinline__ int early_out(const DvrParams& D) { return FAST ? 1 : D.early_out; }
    static __device__ __forceinline__ int blend_mode(const DvrParams& D) { return FAST ? 1 : D.blend_mode; }
    static __device__ __forceinline__ int spp(const DvrParams& D) { return FAST ? 1 : D.spp; }
};
constexpr int kTcComputeThreads = 128;
constexpr int kTcThreads = 160;            // + the MMA warp
constexpr int kTcTileW = 16, kTcTileH = 8;
constexpr int kTcTfBins = 64;              // piecewise TF search accelerator: one start interval per density bin
#ifndef FVSRN_LAGGED_VOTE
#define FVSRN_LAGGED_VOTE 1   // A/B on one B200 (C3): 5.48 ms vs 5.56 ms with a CTA-wide vote every trip
#endif
#ifndef FVSRN_NO_SYNCWARP
#define FVSRN_NO_SYNCWARP 0
#endif
#ifndef FVSRN_TC_ACC16
// 1 (default since r02, applied to the ReLU instantiations - SrnTc::ACC16): fp16 accumulators in TMEM (kind::f16 with D = F16; an element still occupies one 32-bit column), read back
// packed (tcgen05.ld .pack::16b): half the registers per epilogue load and no fp32 -> fp16 packs in the layer epilogues. It is the
// accumulation precision of the reference's own wmma path (fp16 accumulator fragments, renderer_volume_tensorcores.cuh:856-910).
// A/B of full builds on one B200 (r02, tools/ab_full.sh): C3 4.659 -> 4.572 ms, C4 257.7 -> 249.0 ms, C5 1.198 -> 1.176 ms per view;
// C3 frame vs the reference kernel 4.2e-4 -> 5.6e-4 (gate 2e-3); every reference-kernel parity gate of the suite unchanged green.
// 0: fp32 accumulators (the first form of the kernel).
#define FVSRN_TC_ACC16 1
#endif
constexpr int kTcTfFloats = 2560;          // shared-memory TF: up to 204 piecewise intervals x 12 floats, or 256 x float4 texels

// Transfer function staged in shared memory in evaluation-ready form. Piecewise (reference renderer_tf_piecewise.cuh:
// 27-52): one record of three float4 per interval i = (p_i, p_{i+1}, 1/(p_{i+1}-p_i), -), v_i, v_{i+1}-v_i, so that a
// sample costs the interval search plus three LDS.128 and four FFMA. Texture / identity TFs are copied verbatim.
template <int SPEC = 0>
__device__ __forceinline__ void tc_stage_tf(const DvrParams& D, int b, float* stf) {
    using TS = TcSpec<SPEC>;
    if (TS::tf_kind(D) == 0) {
        const float* src = D.tf + (TS::tf_batch(D) == 1 ? 0 : (size_t)b * D.tf_R * 5);
        for (int i = threadIdx.x; i < D.tf_R - 1; i += blockDim.x) {
            const float* v0 = src + i * 5; const float* v1 = v0 + 5;
            const float p0 = __ldg(v0 + 4), p1 = __ldg(v1 + 4);
            float4* rec = reinterpret_cast<float4*>(stf) + 3 * i;
            const float4 a = make_float4(__ldg(v0), __ldg(v0 + 1), __ldg(v0 + 2), __ldg(v0 + 3));
            const float4 e = make_float4(__ldg(v1), __ldg(v1 + 1), __ldg(v1 + 2), __ldg(v1 + 3));
            rec[0] = make_float4(p0, p1, __fdiv_rn(1.0f, p1 - p0), 0.f);
            rec[1] = a;
            rec[2] = make_float4(e.x - a.x, e.y - a.y, e.z - a.z, e.w - a.w);
        }
        // search accelerator: for each of 64 density bins, the interval the reference's linear search (first i with
        // p_{i+1} > d) returns for the bin's lower edge; the search for any d of the bin starts there - same result,
        // usually zero further steps.
        uint8_t* bins = reinterpret_cast<uint8_t*>(stf + 12 * (D.tf_R - 1));
        for (int bin = threadIdx.x; bin < kTcTfBins; bin += blockDim.x) {
            const float d = float(bin) * (1.0f / kTcTfBins);
            int i = 0;
            for (; i < D.tf_R - 2; ++i)
                if (__ldg(src + (i + 1) * 5 + 4) > d) break;
            bins[bin] = uint8_t(i);
        }
    } else {
        const int n = TS::tf_kind(D) == 1 ? D.tf_R * 4 : 2;
        const float* src = D.tf + (TS::tf_batch(D) == 1 ? 0 : (size_t)b * n);
        for (int i = threadIdx.x; i < n; i += blockDim.x) stf[i] = __ldg(src + i);
    }
}
__device__ __forceinline__ float4 tc_tf_piecewise(const float* stf, int R, float d) {
    const float4* rec = reinterpret_cast<const float4*>(stf);
    const uint8_t* bins = reinterpret_cast<const uint8_t*>(stf + 12 * (R - 1));
    int i = bins[min(int(d * float(kTcTfBins)), kTcTfBins - 1)];      // d in [0,1]
    float4 h = rec[3 * i];
    while (i < R - 2 && !(h.y > d)) h = rec[3 * ++i];
    const float4 v0 = rec[3 * i + 1], dv = rec[3 * i + 2];
    const float f = (fminf(fmaxf(d, h.x), h.y) - h.x) * h.z;
    return make_float4(fmaf(f, dv.x, v0.x), fmaf(f, dv.y, v0.y), fmaf(f, dv.z, v0.z), fmaf(f, dv.w, v0.w));
}
// network output -> colour + absorption of one sample; same contract as shade_sample (kernels.cuh), TF from shared memory
template <int SPEC = 0>
__device__ __forceinline__ float4 tc_shade_sample(const NetParams& P, const DvrParams& D, const float* stf, const float (&o)[4]) {
    using TS = TcSpec<SPEC>;
    float4 c = make_float4(0.f, 0.f, 0.f, 0.f);
    if (TS::output_mode(P) < D_OUT_RGBO) {
        const float value = (TS::output_mode(P) == D_OUT_DENSITY) ? sigmoidf_(o[0]) : o[0];
        if (value >= D.dmin) {
            const float d2 = fminf(fmaxf((value - D.dmin) * D.inv_range, 0.f), 1.f);
            if (TS::tf_kind(D) == 0) c = tc_tf_piecewise(stf, D.tf_R, d2);
            else if (TS::tf_kind(D) == 1) c = tf_texture(stf, D.tf_R, d2);
            else c = make_float4(d2 * stf[1], d2 * stf[1], d2 * stf[1], d2 * stf[0]);
            c.w *= D.stepsize;
        }
    } else {    // colour network, SKIP_TRANSFER_FUNCTION path
        if (TS::output_mode(P) == D_OUT_RGBO) c = make_float4(sigmoidf_(o[0]), sigmoidf_(o[1]), sigmoidf_(o[2]), softplusf_(o[3]));
        else c = make_float4(fminf(fmaxf(o[0], 0.f), 1.f), fminf(fmaxf(o[1], 0.f), 1.f), fminf(fmaxf(o[2], 0.f), 1.f), fmaxf(o[3], 0.f));
        c.w *= D.stepsize;
    }
    return c;
}

struct TcShared {           // fixed part of the dynamic shared memory, placed first (48 bytes)
    unsigned long long ready[2];   // slot operands written: one arrival per compute warp
    unsigned long long done[2];    // slot MMA chain complete (tcgen05.commit)
    uint32_t tmem_addr;
    int tile;
    int pad[2];
};

// VARIANT: bit 0 = ReLU hidden activation, bit 1 = NeRF-style Fourier matrix (feature i = axis i%3 times 2^(i/3)),
// bit 2 = SnakeAlt on pre-scaled accumulators (build_tc_blob; exclusive with bit 0)
template <int H16, int GC16, bool DIR, int VARIANT, int SPEC = 0>
struct SrnTc {
    using Core = Srn<H16, GC16, DIR>;
    using TS = TcSpec<SPEC>;
    static constexpr bool RELU = (VARIANT & 1) != 0, NERF = (VARIANT & 2) != 0, SNAKEALT_PRE = (VARIANT & 4) != 0;
    // fp16 accumulators in TMEM for the ReLU instantiations (see FVSRN_TC_ACC16): max() works on the packed pair as it comes out of
    // tcgen05.ld.pack::16b; the SFU activations want fp32 inputs and were 3 % slower with packed accumulators (SnakeAlt 10.6 -> 10.9 ms)
    static constexpr bool ACC16 = (FVSRN_TC_ACC16 != 0) && RELU;
    static constexpr int H = Core::H, K0 = Core::K0, KT0 = K0 / 16, KTH = H / 16;
    static constexpr int NLAST = 16;                       // last layer padded to N = 16 (M=128 needs N % 16 == 0)
    // TMEM columns (32-bit): ones chunk | slot 0: D, A | slot 1: D, A
    static constexpr int ONES_COL = 0, SLOT_COLS = H + K0 / 2, USED_COLS = 8 + 2 * SLOT_COLS;
    static __host__ __device__ constexpr int d_col(int slot) { return 8 + slot * SLOT_COLS; }
    static __host__ __device__ constexpr int a_col(int slot) { return 8 + slot * SLOT_COLS + H; }
    static constexpr int TMEM_COLS = USED_COLS <= 32 ? 32 : USED_COLS <= 64 ? 64 : USED_COLS <= 128 ? 128 : USED_COLS <= 256 ? 256 : 512;
    // weight tiles in the blob (K-major core matrices, chunk-major): bytes
    static __host__ __device__ constexpr int w0_bytes() { return (K0 + 16) * H * 2; }
    static __host__ __device__ constexpr int wh_bytes() { return (H + 16) * H * 2; }
    static __host__ __device__ constexpr int wlast_bytes() { return (H + 16) * NLAST * 2; }

    struct Ctx {
        uint32_t tmem_lane;   // TMEM base address of this thread's lane quadrant (lane field = 32 * (warp % 4))
        uint32_t tmem_base;   // allocation base (lane 0)
        uint32_t sblob;       // shared address of the blob
        uint32_t ready, done; // shared addresses of the mbarrier pairs (slot s at +8 s)
        uint32_t phase;       // parity of the next phase to wait for; both slots advance together
        tc::ActF af;          // (af.act: run-time activation code of the generic instantiations)
    };

    // Layer-0 inputs of the sample owned by this thread (same channel order and fp16 rounding points as
    // Srn::build_row), built for issue-slot economy: Fourier matrix from the constant bank, latent vector through the
    // texture units (4 x tex3D<half4> per 16 channels, hardware trilinear filter = the reference's own fetch,
    // renderer_volume_tensorcores.cuh:573-605), channel 3 = 1 when the network has no time input (bias column).
    // Latent part of the layer-0 row. Issued for both slots of a trip before any other per-sample arithmetic so that the
    // texture latency is covered by the Fourier features of both samples.
    static __device__ __forceinline__ void fetch_latent(const NetParams& P, float px, float py, float pz, uint32_t (&row)[K0 / 2]) {
#pragma unroll
        for (int g = 0; g < GC16; ++g) {
            const float frac = SPEC == 1 ? 0.f : P.groups[g].frac;    // grid-uniform; SPEC 1: static grid, SPEC 2: two keyframes
#ifndef FVSRN_TEX_F16
#define FVSRN_TEX_F16 0   // measured: the f16x2 return path is slower than fp32 return + F2FP (A/B on one B200: 6.05 vs 5.88 ms)
#endif
            if (frac > 0.f || !FVSRN_TEX_F16) {   // temporal lerp of the two keyframes after the fetch, in fp32 (reference :590-604)
                float4 v[4];
#pragma unroll
                for (int q = 0; q < 4; ++q) v[q] = tex3D<float4>(P.tex_a[4 * g + q], px, py, pz);
                if (SPEC == 0 && P.grid_encoding != 0) {        // 8-bit grid (unorm8 textures): decode the filtered texel
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const int ch = 16 * g + 4 * q;
                        v[q].x = decode_grid_value(P.grid_encoding, v[q].x, P.enc_offset_a[ch + 0], P.enc_scale_a[ch + 0]);
                        v[q].y = decode_grid_value(P.grid_encoding, v[q].y, P.enc_offset_a[ch + 1], P.enc_scale_a[ch + 1]);
                        v[q].z = decode_grid_value(P.grid_encoding, v[q].z, P.enc_offset_a[ch + 2], P.enc_scale_a[ch + 2]);
                        v[q].w = decode_grid_value(P.grid_encoding, v[q].w, P.enc_offset_a[ch + 3], P.enc_scale_a[ch + 3]);
                    }
                }
                if (SPEC == 2 || frac > 0.f)
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    float4 w = tex3D<float4>(P.tex_b[4 * g + q], px, py, pz);
                    if (SPEC == 0 && P.grid_encoding != 0) {
                        const int ch = 16 * g + 4 * q;
                        w.x = decode_grid_value(P.grid_encoding, w.x, P.enc_offset_b[ch + 0], P.enc_scale_b[ch + 0]);
                        w.y = decode_grid_value(P.grid_encoding, w.y, P.enc_offset_b[ch + 1], P.enc_scale_b[ch + 1]);
                        w.z = decode_grid_value(P.grid_encoding, w.z, P.enc_offset_b[ch + 2], P.enc_scale_b[ch + 2]);
                        w.w = decode_grid_value(P.grid_encoding, w.w, P.enc_offset_b[ch + 3], P.enc_scale_b[ch + 3]);
                    }
                    v[q].x = fmaf(frac, w.x - v[q].x, v[q].x); v[q].y = fmaf(frac, w.y - v[q].y, v[q].y);
                    v[q].z = fmaf(frac, w.z - v[q].z, v[q].z); v[q].w = fmaf(frac, w.w - v[q].w, v[q].w);
                }
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    row[H / 2 + g * 8 + 2 * q] = tc::pack_h2(v[q].x, v[q].y);
                    row[H / 2 + g * 8 + 2 * q + 1] = tc::pack_h2(v[q].z, v[q].w);
                }
            } else {            // one keyframe: the texture unit returns the filtered texel already packed as 2 x f16x2
#pragma unroll
                for (int q = 0; q < 4; ++q)
                    tc::tex3d_h4(P.tex_a[4 * g + q], px, py, pz, row[H / 2 + g * 8 + 2 * q], row[H / 2 + g * 8 + 2 * q + 1]);
            }
        }
    }

    // The same latent vector through the LSU instead of the texture units, for the world-space batch query (eval_tc_kernel):
    // random positions miss the texture cache (ncu r02a: L1 hit 8.8 %, 16.9 L2 sectors per sample over the four RGBA
    // arrays), while the channel-last copy of the grid keeps all 16 channels of a corner in ONE 32-byte sector: 8 sectors
    // per sample, one 256-bit load (LDG.E.256) per corner = 8 L1 wavefronts per sample. Same addressing and the same 8-bit
    // fractional weights as the texture unit (unnormalised coordinate p*n - 0.5, clamp, weights rounded to 1/256 - oracle
    // latent_group_t); the blend runs in packed fp16, one chain per z plane. Float grids only; byte-encoded grids keep
    // the texture path.
    static __device__ __forceinline__ void ldg256(const void* p, uint32_t (&r)[8]) {
        asm volatile("ld.global.nc.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                     : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]) : "l"(p));
    }
    static __device__ __forceinline__ void fetch_latent_lsu(const NetParams& P, float px, float py, float pz, uint32_t (&row)[K0 / 2]) {
        if (P.grid_encoding != 0) { fetch_latent(P, px, py, pz, row); return; }
        int i0[3], i1[3]; float w1[3];
        const float pp[3] = {px, py, pz};
        const int nn[3] = {P.gx, P.gy, P.gz};
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            const float u = fmaf(pp[a], float(nn[a]), -0.5f);
            const float fl = floorf(u);
            int cell = int(fl);
            int q = __float2int_rn((u - fl) * 256.0f);
            if (q >= 256) { q = 0; cell += 1; }
            i0[a] = min(max(cell, 0), nn[a] - 1);
            i1[a] = min(max(cell + 1, 0), nn[a] - 1);
            w1[a] = float(q) * (1.0f / 256.0f);
        }
#pragma unroll
        for (int g = 0; g < GC16; ++g) {
            const GridGroup& G = P.groups[g];
            const int nsrc = (SPEC == 1 || !(G.frac > 0.f)) ? 1 : 2;     // grid-uniform
            __half2 acc[2][8];
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[0][j] = acc[1][j] = __float2half2_rn(0.f);
#pragma unroll 1
            for (int s = 0; s < nsrc; ++s) {
                const uint4* base = reinterpret_cast<const uint4*>(s == 0 ? G.a : G.b);
                const float ws = (nsrc == 1) ? 1.f : (s == 0 ? 1.f - G.frac : G.frac);
                uint32_t v[8][8];
                __half2 w2[8];
#pragma unroll
                for (int c = 0; c < 8; ++c) {
                    const int zi = (c >> 2) ? i1[2] : i0[2], yi = ((c >> 1) & 1) ? i1[1] : i0[1], xi = (c & 1) ? i1[0] : i0[0];
                    ldg256(base + (size_t)((zi * P.gy + yi) * P.gx + xi) * G.voxel_stride, v[c]);
                    const float wc = ((c >> 2) ? w1[2] : 1.f - w1[2]) * (((c >> 1) & 1) ? w1[1] : 1.f - w1[1]) * ((c & 1) ? w1[0] : 1.f - w1[0]);
                    w2[c] = __float2half2_rn(wc * ws);
                }
#pragma unroll
                for (int c = 0; c < 8; ++c)
#pragma unroll
                    for (int j = 0; j < 8; ++j) acc[c >> 2][j] = __hfma2(w2[c], u2h(v[c][j]), acc[c >> 2][j]);
            }
#pragma unroll
            for (int j = 0; j < 8; ++j) row[H / 2 + g * 8 + j] = h2u(__hadd2(acc[0][j], acc[1][j]));
        }
    }

    // Both samples of a trip at once for the NeRF-style matrix without direction input: the double-angle recurrences of
    // the two samples run in the two lanes of packed-fp32 instructions (FADD2 / FMUL2 / FFMA2, sm_100: one issue slot for
    // both samples; IEEE results, bit-identical to the scalar form in build_row). Everything else as build_row.
    static __device__ __forceinline__ void build_row_pair(const NetParams& P, float px, float py, float pz, float qx, float qy, float qz,
                                                          float dx, float dy, float dz, uint32_t (&row0)[K0 / 2], uint32_t (&row1)[K0 / 2]) {
        if constexpr (NERF && !DIR) {
            constexpr int FOFF = Core::FOFF, NF = Core::NF;
            const __half tv = __float2half_rn(TS::has_time(P) ? P.time_value : 1.f);
            const __half hp[3] = {__float2half_rn(px), __float2half_rn(py), __float2half_rn(pz)};
            const __half hq[3] = {__float2half_rn(qx), __float2half_rn(qy), __float2half_rn(qz)};
            row0[0] = h2u(__halves2half2(hp[0], hp[1])); row0[1] = h2u(__halves2half2(hp[2], tv));
            row1[0] = h2u(__halves2half2(hq[0], hq[1])); row1[1] = h2u(__halves2half2(hq[2], tv));
            float2 cs[NF], sn[NF];      // .x = sample 0, .y = sample 1
#pragma unroll
            for (int a = 0; a < 3; ++a) {
                const uint32_t w = P.fourier2[(a + a * NF) / 2];                       // B[a][a] = entry (a + a*NF)
                const __half b = ((a + a * NF) & 1) ? __high2half(u2h(w)) : __low2half(u2h(w));
                float2 s, c;
                __sincosf(__half2float(__hmul(hp[a], b)), &s.x, &c.x);
                __sincosf(__half2float(__hmul(hq[a], b)), &s.y, &c.y);
                cs[a] = c; sn[a] = s;
#pragma unroll
                for (int i = a + 3; i < NF; i += 3) {
                    const float2 u = __fadd2_rn(s, s);
                    const float2 s2 = __fmul2_rn(u, c);
                    c = __ffma2_rn(make_float2(-u.x, -u.y), s, make_float2(1.0f, 1.0f));
                    s = s2;
                    cs[i] = c; sn[i] = s;
                }
            }
#pragma unroll
            for (int i = 0; i < NF; i += 2) {
                row0[FOFF / 2 + i / 2] = tc::pack_h2(cs[i].x, cs[i + 1].x); row0[FOFF / 2 + NF / 2 + i / 2] = tc::pack_h2(sn[i].x, sn[i + 1].x);
                row1[FOFF / 2 + i / 2] = tc::pack_h2(cs[i].y, cs[i + 1].y); row1[FOFF / 2 + NF / 2 + i / 2] = tc::pack_h2(sn[i].y, sn[i + 1].y);
            }
        } else {
            build_row(P, px, py, pz, dx, dy, dz, row0);
            build_row(P, qx, qy, qz, dx, dy, dz, row1);
        }
    }

    static __device__ __forceinline__ void build_row(const NetParams& P, float px, float py, float pz,
                                                     float dx, float dy, float dz, uint32_t (&row)[K0 / 2]) {
        constexpr int FOFF = Core::FOFF, NF = Core::NF;
        const __half hx = __float2half_rn(px), hy = __float2half_rn(py), hz = __float2half_rn(pz);
        row[0] = h2u(__halves2half2(hx, hy));
        row[1] = h2u(__halves2half2(hz, __float2half_rn(TS::has_time(P) ? P.time_value : 1.f)));
        __half2 d2x, d2y, d2z;
        if (DIR) {
            const __half hdx = __float2half_rn(dx), hdy = __float2half_rn(dy), hdz = __float2half_rn(dz);
            row[2] = h2u(__halves2half2(hdx, hdy));
            row[3] = h2u(__halves2half2(hdz, __float2half_rn(0.f)));
            d2x = __half2half2(hdx); d2y = __half2half2(hdy); d2z = __half2half2(hdz);
        }
        if constexpr (NERF && !DIR) {
            // NeRF-style matrix (network.py:55-63): feature i has the single entry B[i][i%3] = 2^(i/3) * B[i%3][i%3]. The
            // reference's fp16 phase of octave k is then exactly 2^k times the octave-0 phase (power-of-two scaling
            // commutes with rounding), so only the three base phases go through the SFU; higher octaves follow from the
            // double-angle identities in fp32. The error of the seed (sin.approx: ~5e-7 abs) doubles per octave: < 1e-5 after
            // the 4 doublings of the 32-wide network (NF = 14), ~3e-4 after the 9 of the 64-wide one (NF = 30) - there about
            // half an fp16 ulp of the feature, measured end to end: C4 frame vs the reference kernel 6.7e-4 RGBA max-abs
            // (tests/test_gpu_baseline_configs.py C4_relu_nerf_64x6). 6 MUFU instead of 2*NF per sample.
            float cs[NF], sn[NF];
#pragma unroll
            for (int a = 0; a < 3; ++a) {
                const __half h = a == 0 ? hx : (a == 1 ? hy : hz);
                const uint32_t w = P.fourier2[(a + a * NF) / 2];                       // B[a][a] = entry (a + a*NF)
                const __half b = ((a + a * NF) & 1) ? __high2half(u2h(w)) : __low2half(u2h(w));
                float s, c;
                __sincosf(__half2float(__hmul(h, b)), &s, &c);
                cs[a] = c; sn[a] = s;
#pragma unroll
                for (int i = a + 3; i < NF; i += 3) {
                    const float u = s + s;
                    const float s2 = u * c;
                    c = fmaf(-u, s, 1.0f);
                    s = s2;
                    cs[i] = c; sn[i] = s;
                }
            }
#pragma unroll
            for (int i = 0; i < NF; i += 2) {
                row[FOFF / 2 + i / 2] = tc::pack_h2(cs[i], cs[i + 1]);
                row[FOFF / 2 + NF / 2 + i / 2] = tc::pack_h2(sn[i], sn[i + 1]);
            }
        } else {
        const __half2 x2 = __half2half2(hx), y2 = __half2half2(hy), z2 = __half2half2(hz);
    #pragma unroll
            for (int i = 0; i < NF; i += 2) {   // reference :795-807: phase accumulated in fp16 (hmul, hfma, hfma), then cos / sin
                __half2 c = __hmul2(x2, u2h(P.fourier2[(i + 0 * NF) / 2]));
                c = __hfma2(y2, u2h(P.fourier2[(i + 1 * NF) / 2]), c);
                c = __hfma2(z2, u2h(P.fourier2[(i + 2 * NF) / 2]), c);
                if (DIR) {
                    if (P.dir_in_fourier) {
                        c = __hfma2(d2x, u2h(P.fourier2[(i + 3 * NF) / 2]), c);
                        c = __hfma2(d2y, u2h(P.fourier2[(i + 4 * NF) / 2]), c);
                        c = __hfma2(d2z, u2h(P.fourier2[(i + 5 * NF) / 2]), c);
                    }
                }
                const float2 cf = __half22float2(c);
                float s0, c0, s1, c1;
                __sincosf(cf.x, &s0, &c0);
                __sincosf(cf.y, &s1, &c1);
                row[FOFF / 2 + i / 2] = tc::pack_h2(c0, c1);
                row[FOFF / 2 + NF / 2 + i / 2] = tc::pack_h2(s0, s1);
            }
    }
    }

    // MMA warp (all lanes, converged; one elected lane issues): the K steps of one layer of one slot + commit to the
    // slot's completion barrier.
    // The smem descriptor of K step kt differs from step 0 only in the address field of its low word.
    static __device__ __forceinline__ void issue_layer(const Ctx& c, int slot, uint32_t w_addr, int n, int kt_data, bool bias_chunk) {
        const uint32_t idesc = tc::make_idesc(128, n, ACC16);
        const uint32_t kstride = uint32_t(n / 8) * 128u;                    // LBO: distance between the two K chunks of a step
        uint32_t lo = ((w_addr >> 4) & 0x3FFFu) | (((kstride >> 4) & 0x3FFFu) << 16);
        const uint32_t hi = (128u >> 4) | (1u << 14);                       // SBO = 128 B, descriptor version 1
        const uint32_t step = (2u * kstride) >> 4;
        const uint32_t d = c.tmem_base + d_col(slot);
        uint32_t a = c.tmem_base + a_col(slot);
        for (int kt = 0; kt < kt_data; ++kt, lo += step, a += 8) tc::mma_ts2_elect(d, a, lo, hi, idesc, kt > 0);
        if (bias_chunk) tc::mma_ts2_elect(d, c.tmem_base + ONES_COL, lo, hi, idesc, 1u);
        tc::mma_commit_elect(c.done + 8 * slot);
    }

    // per-CTA set-up after TMEM allocation: the constant ones chunk (compute threads)
    static __device__ __forceinline__ void init_tmem(const Ctx& c) {
        uint32_t ones[8] = {0x00003C00u, 0u, 0u, 0u, 0u, 0u, 0u, 0u};   // (half 1.0, half 0), zeros
        tc::tmem_st8(c.tmem_lane + ONES_COL, ones, 0);
        tc::wait_st();
    }

    // compute warp: this warp's operand rows of `slot` are in TMEM -> one arrival on the slot's ready barrier
    static __device__ __forceinline__ void publish(const Ctx& c, int slot) {
        tc::wait_st();
        tc::fence_before_sync();
        if (!FVSRN_NO_SYNCWARP) __syncwarp();    // (elect.sync in the arrive converges the warp as well)
        tc::mbar_arrive_elect(c.ready + 8 * slot);
    }

    // this thread's layer-0 row of one slot -> TMEM A operand
    static __device__ __forceinline__ void store_row(const Ctx& c, int slot, const uint32_t (&row)[K0 / 2]) {
#pragma unroll
        for (int kt = 0; kt < KT0; ++kt) tc::tmem_st8(c.tmem_lane + a_col(slot) + 8 * kt, row, 8 * kt);
        publish(c, slot);
    }

    // epilogue of one slot: accumulator row -> activation -> fp16 A operand of the next layer
    template <int ACT>
    static __device__ __forceinline__ void epilogue_t(const Ctx& c, int slot) {
        if constexpr (ACC16) {
#pragma unroll
        for (int c0 = 0; c0 < H; c0 += 32) {
            uint32_t v[16];
            if (H - c0 >= 32) tc::tmem_ld16_pack(c.tmem_lane + d_col(slot) + c0, v); else tc::tmem_ld8_pack(c.tmem_lane + d_col(slot) + c0, v);
            tc::wait_ld();
#pragma unroll
            for (int j = 0; j < 16; ++j)
                if (H - c0 >= 32 || j < 8) v[j] = tc::act_h2<ACT>(c.af, v[j]);
            if (H - c0 >= 32) tc::tmem_st16(c.tmem_lane + a_col(slot) + c0 / 2, v, 0); else tc::tmem_st8(c.tmem_lane + a_col(slot) + c0 / 2, v, 0);
        }
        } else {
#pragma unroll
        for (int c0 = 0; c0 < H; c0 += 32) {
            uint32_t v[32];
            if (H - c0 >= 32) tc::tmem_ld32(c.tmem_lane + d_col(slot) + c0, v); else tc::tmem_ld16(c.tmem_lane + d_col(slot) + c0, v);
            tc::wait_ld();
            uint32_t pk[16];
#pragma unroll
            for (int j = 0; j < 16; ++j)
                pk[j] = (H - c0 >= 32 || j < 8) ? tc::act_pack<ACT>(c.af, v[2 * j], v[2 * j + 1]) : 0u;
            if (H - c0 >= 32) tc::tmem_st16(c.tmem_lane + a_col(slot) + c0 / 2, pk, 0); else tc::tmem_st8(c.tmem_lane + a_col(slot) + c0 / 2, pk, 0);
        }
        }
    }
    static __device__ __forceinline__ void epilogue(const Ctx& c, int slot, int act) {
        if constexpr (RELU) epilogue_t<D_ACT_RELU>(c, slot);
        else if constexpr (SNAKEALT_PRE) epilogue_t<tc::D_ACT_SNAKEALT_PRE>(c, slot);
        else epilogue_t<-1>(c, slot);       // run-time activation code (c.af.act)
    }

    // Last hidden layer's epilogue of a density network (one output): the 1 x H last layer is a dot product of the
    // activations this thread already holds (fp16x2 FMAs in two interleaved chains, summed in fp32), as the reference
    // does it on the CUDA cores (renderer_volume_tensorcores.cuh:1138-1143) - no further tensor-core round trip.
    template <int ACT>
    static __device__ __forceinline__ float epilogue_dot_t(const NetParams& P, const Ctx& c, int slot) {
        // two interleaved fp16x2 chains per block of 32 channels, each block summed into fp32: wide networks (H = 96, 128) do
        // not accumulate more than 8 products per fp16 lane (W128x2 frame vs the reference kernel: 2.9e-3 with one chain pair
        // over all 128 channels)
        float sum = P.blast;
#pragma unroll
        for (int c0 = 0; c0 < H; c0 += 32) {
            __half2 a0 = __float2half2_rn(0.f), a1 = a0;
            if constexpr (ACC16) {
                uint32_t v[16];
                if (H - c0 >= 32) tc::tmem_ld16_pack(c.tmem_lane + d_col(slot) + c0, v); else tc::tmem_ld8_pack(c.tmem_lane + d_col(slot) + c0, v);
                tc::wait_ld();
#pragma unroll
                for (int j = 0; j < 16; j += 2) {
                    if (H - c0 >= 32 || j < 8) {
                        const uint32_t p0 = tc::act_h2<ACT>(c.af, v[j]), p1 = tc::act_h2<ACT>(c.af, v[j + 1]);
                        a0 = __hfma2(u2h(p0), u2h(P.wlast2[c0 / 2 + j]), a0);
                        a1 = __hfma2(u2h(p1), u2h(P.wlast2[c0 / 2 + j + 1]), a1);
                    }
                }
            } else {
                uint32_t v[32];
                if (H - c0 >= 32) tc::tmem_ld32(c.tmem_lane + d_col(slot) + c0, v); else tc::tmem_ld16(c.tmem_lane + d_col(slot) + c0, v);
                tc::wait_ld();
#pragma unroll
                for (int j = 0; j < 16; j += 2) {
                    if (H - c0 >= 32 || j < 8) {
                        const uint32_t p0 = tc::act_pack<ACT>(c.af, v[2 * j], v[2 * j + 1]), p1 = tc::act_pack<ACT>(c.af, v[2 * j + 2], v[2 * j + 3]);
                        a0 = __hfma2(u2h(p0), u2h(P.wlast2[c0 / 2 + j]), a0);
                        a1 = __hfma2(u2h(p1), u2h(P.wlast2[c0 / 2 + j + 1]), a1);
                    }
                }
            }
            const float2 f0 = __half22float2(a0), f1 = __half22float2(a1);
            sum += (f0.x + f0.y) + (f1.x + f1.y);
        }
        return sum;
    }
    static __device__ __forceinline__ float epilogue_dot(const NetParams& P, const Ctx& c, int slot, int act) {
        if constexpr (RELU) return epilogue_dot_t<D_ACT_RELU>(P, c, slot);
        else if constexpr (SNAKEALT_PRE) return epilogue_dot_t<tc::D_ACT_SNAKEALT_PRE>(P, c, slot);
        else return epilogue_dot_t<-1>(P, c, slot);
    }

    // Compute warps: run the layer chain for the 2 x 128 samples whose layer-0 rows were stored with store_row().
    // A warp that is not `warp_active` (all its rays finished) keeps the barrier protocol going but skips the arithmetic.
    static __device__ __forceinline__ void mlp_compute(const NetParams& P, Ctx& c, bool warp_active, float (&out0)[4], float (&out1)[4]) {
        const int act = P.activation;
        const bool scalar_last = TS::output_mode(P) < D_OUT_RGBO;    // density networks: last layer in registers
        // hidden-to-hidden hand-overs: accumulator -> activation -> A operand of the next layer
#pragma unroll 1
        for (int l = scalar_last ? 1 : 0; l <= P.num_hidden; ++l) {
#pragma unroll
            for (int slot = 0; slot < 2; ++slot) {
                tc::mbar_wait(c.done + 8 * slot, c.phase);
                tc::fence_after_sync();
                if (warp_active) epilogue(c, slot, act);
                publish(c, slot);
            }
            c.phase ^= 1u;
        }
        if (scalar_last) {   // last hidden layer: activation + 1 x H dot product in registers
#pragma unroll
            for (int slot = 0; slot < 2; ++slot) {
                float (&out)[4] = slot == 0 ? out0 : out1;
                tc::mbar_wait(c.done + 8 * slot, c.phase);
                tc::fence_after_sync();
                out[0] = warp_active ? epilogue_dot(P, c, slot, act) : 0.f;
                out[1] = out[2] = out[3] = 0.f;
            }
            c.phase ^= 1u;
        }
        if (scalar_last) { tc::fence_before_sync(); return; }
#pragma unroll
        for (int slot = 0; slot < 2; ++slot) {
            float (&out)[4] = slot == 0 ? out0 : out1;
            tc::mbar_wait(c.done + 8 * slot, c.phase);
            tc::fence_after_sync();
            if (TS::output_mode(P) >= D_OUT_RGBO) {
                uint32_t o[4];
                tc::tmem_ld4(c.tmem_lane + d_col(slot), o);
                tc::wait_ld();
                if constexpr (ACC16) {
#pragma unroll
                    for (int k = 0; k < 4; ++k) out[k] = __half2float(__ushort_as_half((unsigned short)(o[k] & 0xffffu)));
                } else {
                    out[0] = __uint_as_float(o[0]); out[1] = __uint_as_float(o[1]); out[2] = __uint_as_float(o[2]); out[3] = __uint_as_float(o[3]);
                }
            } else {
                uint32_t o;
                tc::tmem_ld1(c.tmem_lane + d_col(slot), o);
                tc::wait_ld();
                if constexpr (ACC16) { out[0] = __half2float(__ushort_as_half((unsigned short)(o & 0xffffu))); out[1] = out[2] = out[3] = 0.f; }
                else { out[0] = __uint_as_float(o); out[1] = out[2] = out[3] = 0.f; }
            }
        }
        c.phase ^= 1u;
        // the next trip's tcgen05.st / mma may only start after this thread has read its outputs
        tc::fence_before_sync();
    }

    // MMA thread: the matching issue schedule of one trip. `c.phase` here is the parity of the ready barriers.
    static __device__ __forceinline__ void mlp_issue(const NetParams& P, Ctx& c) {
        const int last = (TS::output_mode(P) < D_OUT_RGBO) ? P.num_hidden - 1 : P.num_hidden;   // see epilogue_dot
#pragma unroll 1
        for (int l = -1; l <= last; ++l) {
#pragma unroll
            for (int slot = 0; slot < 2; ++slot) {
                tc::mbar_wait(c.ready + 8 * slot, c.phase);
                tc::fence_after_sync();
                if (l < 0) issue_layer(c, slot, c.sblob + P.off_w0, H, KT0, TS::has_time(P) != 0);
                else if (l < P.num_hidden) issue_layer(c, slot, c.sblob + P.off_wh + l * wh_bytes(), H, KTH, true);
                else issue_layer(c, slot, c.sblob + P.off_wlast, NLAST, KTH, true);
            }
            c.phase ^= 1u;
        }
    }
};

// common prologue: blob -> smem, mbarriers, TMEM allocation (by the MMA warp)
template <class Net>
__device__ __forceinline__ typename Net::Ctx tc_prologue(const NetParams& P, TcShared* sh, uint4* sblob) {
    load_blob(P, sblob);
    const int warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) {
        sh->pad[0] = sh->pad[1] = -1;    // trip-vote flags (FVSRN_LAGGED_VOTE)
        tc::mbar_init(smem_u32(&sh->ready[0]), 4); tc::mbar_init(smem_u32(&sh->ready[1]), 4);
        tc::mbar_init(smem_u32(&sh->done[0]), 1); tc::mbar_init(smem_u32(&sh->done[1]), 1);
    }
    if (warp == 4) tc::tmem_alloc(smem_u32(&sh->tmem_addr), Net::TMEM_COLS);
    // make the generic-proxy writes of the weights visible to the tensor-core (async) proxy
    asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    typename Net::Ctx c;
    c.tmem_base = sh->tmem_addr;
    c.tmem_lane = c.tmem_base + (uint32_t((warp & 3) * 32) << 16);
    asm volatile("" : "+r"(c.tmem_lane));   // opaque: kept in a register instead of being re-derived from %tid at every TMEM access
    c.sblob = smem_u32(sblob);
    c.ready = smem_u32(&sh->ready[0]);
#ifndef FVSRN_OPAQUE_BAR
#define FVSRN_OPAQUE_BAR 1
#endif
    if (FVSRN_OPAQUE_BAR) asm volatile("" : "+r"(c.ready));   // opaque: not re-derived (S2R SR_CgaCtaId + LEA) at every wait
    c.done = c.ready + 16;
    c.phase = 0;
    c.af = tc::make_actf(P.activation, P.act_param);
    if (warp < 4) Net::init_tmem(c);
    return c;
}
template <class Net>
__device__ __forceinline__ void tc_epilogue(const typename Net::Ctx& c) {
    tc::fence_before_sync();
    __syncthreads();
    if ((threadIdx.x >> 5) == 4) tc::tmem_dealloc(c.tmem_base, Net::TMEM_COLS);
}

template <int H16, int GC16, bool DIR, int VARIANT, int SPEC = 0>
// Register budget: 4 CTAs/SM at hidden <= 32 (96 registers); 2 CTAs/SM at hidden 48/64, which needs <= 168 registers
// because a warp's registers must fit one 16384-entry register-file partition three times over (see
// tc_blocks_by_regs in launch.cuh) - expressed as (192 threads, 2 blocks), the launch still uses kTcThreads.
__global__ void __launch_bounds__(H16 <= 2 ? kTcThreads : 192, H16 <= 2 ? 4 : (H16 <= 4 ? 2 : 1))
dvr_tc_kernel(const __grid_constant__ NetParams P, const __grid_constant__ DvrParams D) {
    using Net = SrnTc<H16, GC16, DIR, VARIANT, SPEC>;
    using TS = TcSpec<SPEC>;
    using Core = typename Net::Core;
    extern __shared__ uint4 smem[];
    TcShared* sh = reinterpret_cast<TcShared*>(smem);
    uint4* sblob = smem + 3;
    const int blob16 = (P.blob_bytes + 15) / 16;
    float* stf = reinterpret_cast<float*>(sblob + blob16);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const bool is_mma_warp = warp == 4;

    typename Net::Ctx ctx = tc_prologue<Net>(P, sh, sblob);
    if (TS::tf_batch(D) == 1) tc_stage_tf<SPEC>(D, 0, stf);
    __syncthreads();

    const float3 bmin = make_float3(P.box_min[0], P.box_min[1], P.box_min[2]);
    const float3 bsize = make_float3(P.box_size[0], P.box_size[1], P.box_size[2]);
    const float3 binv = make_float3(__fdiv_rn(1.f, bsize.x), __fdiv_rn(1.f, bsize.y), __fdiv_rn(1.f, bsize.z));
    // pixel of this thread inside the CTA tile: warps 2x2, lanes 8x4
    const int lx = (warp & 1) * 8 + (lane & 7), ly = ((warp >> 1) & 1) * 4 + (lane >> 3);

    unsigned long long n_valid = 0, n_exec = 0;
    int trip_id = 0;     // trips executed by this CTA (same count in every warp)
    for (;;) {
        if (threadIdx.x == 0) sh->tile = int(atomicAdd(D.tile_counter, 1u));
        __syncthreads();
        const int tile = sh->tile;
        __syncthreads();
        if (tile >= D.num_tiles) break;
        const int tiles_per_img = D.tiles_x * D.tiles_y;
        const int bl = tile / tiles_per_img;
        const int trem = tile - bl * tiles_per_img;
        const int ty = trem / D.tiles_x, tx = trem - ty * D.tiles_x;
        const int b = D.b0 + bl;
        if (TS::tf_batch(D) != 1) {      // per-view transfer function: staged for the view this tile belongs to
            tc_stage_tf<SPEC>(D, b, stf);
            __syncthreads();
        }

        if (is_mma_warp) {
            // ---------------------------------------------------------------- MMA warp: one elected thread issues
            for (int s = 0; s < TS::spp(D); ++s) {
                if (FVSRN_LAGGED_VOTE) {
                    // one CTA vote per march; afterwards trip i+1 runs iff some warp had work in trip i (flag written
                    // before that warp's first arrival of trip i, read after this warp's last ready-wait of trip i)
                    bool cont = __syncthreads_or(0);
                    while (cont) {
                        Net::mlp_issue(P, ctx);
                        cont = *reinterpret_cast<volatile int*>(&sh->pad[trip_id & 1]) == trip_id;
                        ++trip_id;
                    }
                } else
                for (;;) {
                    if (!__syncthreads_or(0)) break;
                    Net::mlp_issue(P, ctx);
                }
            }
            continue;
        }
        // -------------------------------------------------------------------- compute warps
        const int x = D.x0 + tx * kTcTileW + lx;
        const int y = D.y0 + ty * kTcTileH + ly;
        const bool in_img = (x < D.x0 + D.tw) && (y < D.y0 + D.th);
        const CamFrame cf = load_camera(D, b);

        float4 pix_color = make_float4(0.f, 0.f, 0.f, 0.f);
        float pix_depth = 0.f;
        int last_steps = 0;
        for (int s = 0; s < TS::spp(D); ++s) {   // CTA-uniform trip count
            const Ray r = make_ray(D, cf, x, y, s, bmin, bsize, TS::spp(D));
            const float tmax = in_img ? r.tmax : -1.f;
            float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
            float acc_depth = 0.f;
            int steps = 0;
            // lock-step march of the 128 rays, reference renderer_ray_evaluation_stepping_dvr.cuh:84-154. Two consecutive
            // steps are evaluated per trip (slot 0 = step i, slot 1 = step i+1, the second speculatively: whether it is
            // needed is only known after step i has been composited); samples are blended in order, so the result is
            // the sequential one.
            bool cont = true;
            for (int i = 0;; i += 2) {
                const float t0 = __fmaf_rn(float(i), D.stepsize, r.tmin);
                const float t1 = __fmaf_rn(float(i + 1), D.stepsize, r.tmin);
                const bool alive = !TS::early_out(D) || (acc.w < D.alpha_early_out);
                const bool valid0 = (t0 <= tmax) && alive;
                const bool in1 = (t1 <= tmax) && alive;
                if (FVSRN_LAGGED_VOTE) {
                    if (i == 0) cont = __syncthreads_or(valid0 ? 1 : 0);
                    if (!cont) break;
                } else if (!__syncthreads_or(valid0 ? 1 : 0)) break;
                // a warp whose 32 rays are all finished keeps the barrier protocol going but skips the arithmetic
                const bool warp_active = __any_sync(0xffffffffu, valid0);
                if (FVSRN_LAGGED_VOTE && warp_active && lane == 0) *reinterpret_cast<volatile int*>(&sh->pad[trip_id & 1]) = trip_id;
                if (warp_active) {
                    float px = 0.f, py = 0.f, pz = 0.f, qx = 0.f, qy = 0.f, qz = 0.f;
                    if (valid0) {
                        px = __fmul_rn(__fsub_rn(__fmaf_rn(r.dir.x, t0, r.eye.x), bmin.x), binv.x);
                        py = __fmul_rn(__fsub_rn(__fmaf_rn(r.dir.y, t0, r.eye.y), bmin.y), binv.y);
                        pz = __fmul_rn(__fsub_rn(__fmaf_rn(r.dir.z, t0, r.eye.z), bmin.z), binv.z);
                    }
                    if (in1) {
                        qx = __fmul_rn(__fsub_rn(__fmaf_rn(r.dir.x, t1, r.eye.x), bmin.x), binv.x);
                        qy = __fmul_rn(__fsub_rn(__fmaf_rn(r.dir.y, t1, r.eye.y), bmin.y), binv.y);
                        qz = __fmul_rn(__fsub_rn(__fmaf_rn(r.dir.z, t1, r.eye.z), bmin.z), binv.z);
                    }
                    uint32_t row0[Core::K0 / 2], row1[Core::K0 / 2];
#ifndef FVSRN_TEX_EARLY
#define FVSRN_TEX_EARLY 1
#endif
                    Net::fetch_latent(P, px, py, pz, row0);
                    if (FVSRN_TEX_EARLY) Net::fetch_latent(P, qx, qy, qz, row1);
#ifndef FVSRN_PAIR_FOURIER
#define FVSRN_PAIR_FOURIER 0   // measured (A/B on one B200, C3): 5.59 ms with the packed-fp32 pair form vs 5.55 ms scalar - 33 fewer issue slots per trip, but slot 0 is published later
#endif
                    if (FVSRN_PAIR_FOURIER && FVSRN_TEX_EARLY) {
                        Net::build_row_pair(P, px, py, pz, qx, qy, qz, r.dir.x, r.dir.y, r.dir.z, row0, row1);
                        Net::store_row(ctx, 0, row0);
                        Net::store_row(ctx, 1, row1);
                    } else {
                        Net::build_row(P, px, py, pz, r.dir.x, r.dir.y, r.dir.z, row0);
                        Net::store_row(ctx, 0, row0);
                        if (!FVSRN_TEX_EARLY) Net::fetch_latent(P, qx, qy, qz, row1);
                        Net::build_row(P, qx, qy, qz, r.dir.x, r.dir.y, r.dir.z, row1);
                        Net::store_row(ctx, 1, row1);
                    }
                } else {
                    Net::publish(ctx, 0);
                    Net::publish(ctx, 1);
                }
                float o0[4], o1[4];
                Net::mlp_compute(P, ctx, warp_active, o0, o1);
                if (FVSRN_LAGGED_VOTE) {    // every warp's flag store of this trip precedes the MMAs this warp just waited for
                    cont = *reinterpret_cast<volatile int*>(&sh->pad[trip_id & 1]) == trip_id;
                    ++trip_id;
                }
                if (lane == 0) n_exec += 64;
                // both samples are shaded first (independent work, interleaved by the compiler), then composited in order
                const float4 c0 = tc_shade_sample<SPEC>(P, D, stf, o0), c1 = tc_shade_sample<SPEC>(P, D, stf, o1);
                steps += valid0 ? 1 : 0;
                blend_sample_pred(D, TS::blend_mode(D), valid0, c0, t0, acc, acc_depth);
                const bool valid1 = in1 && (!TS::early_out(D) || (acc.w < D.alpha_early_out));
                steps += valid1 ? 1 : 0;
                blend_sample_pred(D, TS::blend_mode(D), valid1, c1, t1, acc, acc_depth);
            }
            n_valid += steps;
            last_steps = steps;
            pix_color.x += acc.x; pix_color.y += acc.y; pix_color.z += acc.z; pix_color.w += acc.w;
            pix_depth += acc_depth * acc.w;
        }
        if (in_img) write_pixel(D, b, x, y, pix_color, pix_depth, last_steps, TS::spp(D));
    }
    if (D.counters && !is_mma_warp) {
        for (int off = 16; off > 0; off >>= 1) n_valid += __shfl_xor_sync(0xffffffffu, n_valid, off);
        if (lane == 0) { atomicAdd(D.counters + 0, n_valid); atomicAdd(D.counters + 1, n_exec); }
    }
    tc_epilogue<Net>(ctx);
}

template <int H16, int GC16, bool DIR, int VARIANT>
__global__ void __launch_bounds__(H16 <= 2 ? kTcThreads : 192, H16 <= 2 ? 4 : (H16 <= 4 ? 2 : 1))
eval_tc_kernel(const __grid_constant__ NetParams P, const __grid_constant__ EvalParams E) {
    using Net = SrnTc<H16, GC16, DIR, VARIANT>;
    using TS = TcSpec<0>;
    using Core = typename Net::Core;
    extern __shared__ uint4 smem[];
    TcShared* sh = reinterpret_cast<TcShared*>(smem);
    uint4* sblob = smem + 3;
    typename Net::Ctx ctx = tc_prologue<Net>(P, sh, sblob);
    const long long n256 = (E.n + 255) / 256;
    if ((threadIdx.x >> 5) == 4) {
        for (long long w = blockIdx.x; w < n256; w += gridDim.x) Net::mlp_issue(P, ctx);
    } else {
        for (long long w = blockIdx.x; w < n256; w += gridDim.x) {
            long long idx[2]; bool valid[2];
            float pos[2][3], dir[2][3];
#pragma unroll
            for (int slot = 0; slot < 2; ++slot) {
                const long long i = w * 256 + slot * 128 + threadIdx.x;
                idx[slot] = i; valid[slot] = i < E.n;
#pragma unroll
                for (int a = 0; a < 3; ++a) {
                    pos[slot][a] = valid[slot] ? __ldg(E.pos + 3 * i + a) : 0.f;
                    dir[slot][a] = (DIR && E.dir && valid[slot]) ? __ldg(E.dir + 3 * i + a) : 0.f;
                }
            }
#ifndef FVSRN_EVAL_ORDER
#define FVSRN_EVAL_ORDER 0   // measured on C2 (random positions): 0 = Fourier first, then fetch: 0.844 ms; 1 = both fetches first: 0.861 ms
#endif
#ifndef FVSRN_EVAL_LSU
// latent fetch of the batch query: 1 = channel-last 256-bit loads (8 L2 sectors per sample), 0 = texture units (16.9).
// A/B on one B200 (C2, 2^24 random positions, r02h): 0.890 ms through the texture units, 0.837 with 16 LDG.128 per sample,
// 0.660 with 8 LDG.256; with two keyframes 1.568 -> 1.098 ms. A lane-pair variant (even lane x0 corner, odd lane x1 corner of
// the same sample: 5 instead of 8 L1 wavefronts per sample, partial sums exchanged by SHFL) measured 0.758 - it spills at 96 registers.
#define FVSRN_EVAL_LSU 1
#endif
            uint32_t row0[Core::K0 / 2], row1[Core::K0 / 2];
            auto fetch = [&](int slot, uint32_t (&row)[Core::K0 / 2]) {
                if (FVSRN_EVAL_LSU) Net::fetch_latent_lsu(P, pos[slot][0], pos[slot][1], pos[slot][2], row);
                else Net::fetch_latent(P, pos[slot][0], pos[slot][1], pos[slot][2], row);
            };
            if (FVSRN_EVAL_ORDER == 1) {        // both slots' latent fetches in flight before any other arithmetic
                fetch(0, row0);
                fetch(1, row1);
            }
            Net::build_row(P, pos[0][0], pos[0][1], pos[0][2], dir[0][0], dir[0][1], dir[0][2], row0);
            if (FVSRN_EVAL_ORDER == 0) fetch(0, row0);
            Net::store_row(ctx, 0, row0);
            Net::build_row(P, pos[1][0], pos[1][1], pos[1][2], dir[1][0], dir[1][1], dir[1][2], row1);
            if (FVSRN_EVAL_ORDER == 0) fetch(1, row1);
            Net::store_row(ctx, 1, row1);
            float o0[4], o1[4];
            Net::mlp_compute(P, ctx, true, o0, o1);
#pragma unroll
            for (int slot = 0; slot < 2; ++slot) {
                if (!valid[slot]) continue;
                const float (&o)[4] = slot == 0 ? o0 : o1;
                const long long i = idx[slot];
                // output parametrisation, reference :1054-1072, 1149-1158 (value, not valueNoClamping)
                if (TS::output_mode(P) == D_OUT_DENSITY) E.out[i] = sigmoidf_(o[0]);
                else if (TS::output_mode(P) == D_OUT_DENSITY_DIRECT) E.out[i] = o[0];
                else {
                    float4 c;
                    if (TS::output_mode(P) == D_OUT_RGBO) c = make_float4(sigmoidf_(o[0]), sigmoidf_(o[1]), sigmoidf_(o[2]), softplusf_(o[3]));
                    else c = make_float4(fminf(fmaxf(o[0], 0.f), 1.f), fminf(fmaxf(o[1], 0.f), 1.f), fminf(fmaxf(o[2], 0.f), 1.f), fmaxf(o[3], 0.f));
                    reinterpret_cast<float4*>(E.out)[i] = c;
                }
            }
        }
    }
    tc_epilogue<Net>(ctx);
}

} // namespace fvsrn
