// Reference-exact mode (FVSRN_KERNEL_REFERENCE_EXACT): the DVR frame with the arithmetic of the reference's compiled
// kernel, operation for operation, so that images AND early-termination step counts are bit-identical to
// ImageEvaluatorSimpleKernel + RayEvaluationSteppingDvr + VolumeInterpolationTensorcores run on the same GPU
// (north_star: "early-termination step counts bit-exact against the reference's tensor-core kernel when it is run
// with the same precision"). It is the validation mode, not the fast path.
//
// What "same arithmetic" means here (reference file:line; the operation order is that of the reference kernel as NVRTC
// compiles it with --use_fast_math, kernel_loader.cpp:229-241):
//   * ray setup, stepping, position: rf:: operations (flush-to-zero, div.approx, rsqrt.approx, sin/cos.approx), kernels.cuh
//   * p = (pos - boxMin) / boxSize with div.approx (renderer_volume_tensorcores.cuh:746); fp16 inputs by cvt.rn (:768-771)
//   * Fourier features: fp16 phase chain hmul, hfma, hfma, then hcos / hsin of cuda_fp16.hpp (:795-807)
//   * latent vector: tex3D<float4> on FP32 textures, both keyframes always fetched, v = vA + frac(t)*(vB - vA) as
//     sub + fma, then cvt.rn.f16 (:573-605)
//   * layers: HMMA m16n8k16 with fp16 accumulators, accumulator preloaded with the bias, K consumed 16 at a time,
//     Fourier block first, then the latent chunk (:839-910, 956-995: wmma m16n16k16 = two HMMA.16816.F16 per op)
//   * activations on the fp16 accumulator elements with the reference's formulas for __CUDA_ARCH__ < 800 - the kernel is
//     built for compute_75 (renderer_activations.cuh:39-341)
//   * last layer: one sequential fp16 fma chain over the hidden channels, starting from the bias (:1138-1143)
//   * density range, piecewise / texture TF, Beer-Lambert / alpha blending and the output normalisation in the
//     reference's fp32 operation order (renderer_ray_evaluation_stepping_dvr.cuh:84-154, renderer_tf_piecewise.cuh:27-62,
//     renderer_tf_texture.cuh:46-56, renderer_blending.cuh:36-51, renderer_image_evaluator_simple.cuh:96-124)
// Scope: density-type networks (density[:direct], densitygrad[:direct|:cubic], densitycurvature[:direct]), Fourier
// features, one 16-channel latent grid group (static, time keyframes or ensemble) for every supported width, two groups and / or the ray direction as input (also inside the Fourier
// features) for the 32- and 64-wide networks, optional time input; normals
// predicted by the network or by finite differences, BRDFLambert shading - the shapes the reference kernels
// in oracle/_ref are built for.
#pragma once
#include "kernels.cuh"

namespace fvsrn {

// hcos / hsin of cuda_fp16.hpp (range reduction + polynomial + the special-case fix-ups, ~40 instructions each) as out-of-line
// functions: inlined at each of their ~100 call sites they made the kernels 150 KB of code and the march instruction-fetch bound
// (ncu r02n: stall no_instruction 14.7 per issue). Same instructions, same results.
static __device__ __noinline__ __half exact_hcos(__half x) { return hcos(x); }
static __device__ __noinline__ __half exact_hsin(__half x) { return hsin(x); }
static __device__ __noinline__ float exact_erfinv(float x) { return erfinvf(x); }     // byte-gaussian grids: 32 call sites per latent group

template <int H16, int GC16 = 1, bool DIR = false>
struct SrnExact {
    using Core = Srn<H16, GC16, DIR>;
    static constexpr int H = Core::H, K0 = Core::K0, NF = Core::NF, FOFF = Core::FOFF, NT = Core::NT, KT0 = Core::KT0, KTH = Core::KTH;
    static constexpr int RS_CHUNKS = Core::RS_CHUNKS, ROW_CHUNKS = Core::ROW_CHUNKS;

    struct Act {        // per-network constants of the hidden activation, as the reference forms them per call
        __half p;       // Sine / Snake: half(param)
        __half q;       // Snake: half(1/param) (fast-math division); SnakeAlt: half(2*param)
        __half one;
    };
    static __device__ __forceinline__ Act make_act(int act, float param) {
        Act a;
        a.p = __float2half(param);
        a.q = (act == D_ACT_SNAKE) ? __float2half(rf::div(1.0f, param)) : __float2half(rf::mul(param, 2.0f));
        a.one = __float2half(1.0f);
        return a;
    }
    static __device__ __forceinline__ __half act1(int act, const Act& a, __half v) {
        if (act == D_ACT_RELU) return __float2half(rf::fmax(0.f, __half2float(v)));          // :46
        if (act == D_ACT_SINE) return exact_hsin(__hmul(v, a.p));                                   // :103
        if (act == D_ACT_SNAKE) { const __half s = exact_hsin(__hmul(a.p, v)); return __hadd(v, __hmul(a.q, __hmul(s, s))); }   // :266-269
        if (act == D_ACT_SNAKEALT) { const __half x0 = exact_hcos(__hmul(a.q, v)); return __hdiv(__hsub(__hadd(v, a.one), x0), a.q); }  // :333-337
        return v;
    }
    static __device__ __forceinline__ void activate(uint32_t (&r)[2][NT][2], int act, const Act& a) {
#pragma unroll
        for (int mt = 0; mt < 2; ++mt)
#pragma unroll
            for (int nt = 0; nt < NT; ++nt)
#pragma unroll
                for (int j = 0; j < 2; ++j) {
                    const __half2 v = u2h(r[mt][nt][j]);
                    r[mt][nt][j] = h2u(__halves2half2(act1(act, a, __low2half(v)), act1(act, a, __high2half(v))));
                }
    }

    static __device__ __forceinline__ void copy_frag(uint32_t (&dst)[2][NT][2], const uint32_t (&src)[2][NT][2]) {
#pragma unroll
        for (int mt = 0; mt < 2; ++mt)
#pragma unroll
            for (int nt = 0; nt < NT; ++nt) { dst[mt][nt][0] = src[mt][nt][0]; dst[mt][nt][1] = src[mt][nt][1]; }
    }
    // accumulator fragments <-> one row per sample in the warp's staging tile: register j of tile (mt, nt) holds sample
    // mt*16 + lane/4 + 8*j, channels nt*8 + 2*(lane%4) + {0,1}
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
    static __device__ __forceinline__ void rows_to_frag(const uint4* stage, uint32_t (&f)[2][NT][2], int lane) {
        const uint32_t* st32 = reinterpret_cast<const uint32_t*>(stage);
#pragma unroll
        for (int mt = 0; mt < 2; ++mt)
#pragma unroll
            for (int nt = 0; nt < NT; ++nt)
#pragma unroll
                for (int j = 0; j < 2; ++j)
                    f[mt][nt][j] = st32[(mt * 16 + (lane >> 2) + 8 * j) * (RS_CHUNKS * 4) + nt * 4 + (lane & 3)];
    }

    // layer-0 inputs of the sample owned by this lane: [px,py,pz,0, cos x NF, sin x NF, latent x 16]
    static __device__ __forceinline__ void build_row(const NetParams& P, float px, float py, float pz, float3 dir, float tfrac, uint32_t (&row)[K0 / 2]) {
        const __half vx = __float2half(px), vy = __float2half(py), vz = __float2half(pz);
        row[0] = h2u(__halves2half2(vx, vy));
        // input channel 3: the keyframe time when PASS_TIME_TO_NETWORK (cLatentGridInterpolation[0].x, :777-779), else 0
        row[1] = h2u(__halves2half2(vz, __float2half(P.has_time ? P.time_value : 0.f)));
        const __half dx = __float2half(dir.x), dy = __float2half(dir.y), dz = __float2half(dir.z);
        if constexpr (DIR) {            // USE_DIRECTION >= 1: input channels 4..7 = direction, 0 (:784-792)
            row[2] = h2u(__halves2half2(dx, dy));
            row[3] = h2u(__halves2half2(dz, __float2half(0.f)));
        }
        const __half* F = reinterpret_cast<const __half*>(P.fourier2);
        __half cs[NF], sn[NF];
#pragma unroll
        for (int i = 0; i < NF; ++i) {
            __half c = __hmul(vx, F[i]);
            c = __hfma(vy, F[i + NF], c);
            c = __hfma(vz, F[i + 2 * NF], c);
            if (DIR && P.dir_in_fourier) {      // USE_DIRECTION == 2 (:800-804)
                c = __hfma(dx, F[i + 3 * NF], c);
                c = __hfma(dy, F[i + 4 * NF], c);
                c = __hfma(dz, F[i + 5 * NF], c);
            }
            cs[i] = exact_hcos(c); sn[i] = exact_hsin(c);
        }
#pragma unroll
        for (int i = 0; i < NF; i += 2) {
            row[FOFF / 2 + i / 2] = h2u(__halves2half2(cs[i], cs[i + 1]));
            row[FOFF / 2 + NF / 2 + i / 2] = h2u(__halves2half2(sn[i], sn[i + 1]));
        }
#pragma unroll
        for (int q = 0; q < 4 * GC16; ++q) {
            float4 a = tex3D<float4>(P.tex_a[q], px, py, pz);
            float4 b = tex3D<float4>(P.tex_b[q], px, py, pz);
            if (P.grid_encoding != 0) {     // EncodeGridValue<Encoding, false> on BOTH keyframes: the reference decodes B with A's constants (:586-587)
                float* pa = &a.x; float* pb = &b.x;
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    float ta = pa[k], tb = pb[k];
                    if (P.grid_encoding == 2) {
                        ta = rf::mul(1.4142135623730951f, exact_erfinv(rf::mul(rf::add(ta, -0.5f), 2.0f - 1e-4f)));
                        tb = rf::mul(1.4142135623730951f, exact_erfinv(rf::mul(rf::add(tb, -0.5f), 2.0f - 1e-4f)));
                    }
                    pa[k] = rf::fma(ta, P.enc_scale_a[4 * q + k], P.enc_offset_a[4 * q + k]);
                    pb[k] = rf::fma(tb, P.enc_scale_a[4 * q + k], P.enc_offset_a[4 * q + k]);
                }
            }
            const float x = rf::fma(rf::sub(b.x, a.x), tfrac, a.x), y = rf::fma(rf::sub(b.y, a.y), tfrac, a.y);
            const float z = rf::fma(rf::sub(b.z, a.z), tfrac, a.z), w = rf::fma(rf::sub(b.w, a.w), tfrac, a.w);
            row[H / 2 + 2 * q] = h2u(__halves2half2(__float2half(x), __float2half(y)));
            row[H / 2 + 2 * q + 1] = h2u(__halves2half2(__float2half(z), __float2half(w)));
        }
    }

    // network output (raw last-layer values, fp16) of the sample owned by this lane; all 32 lanes call. four = the
    // 4 / 6-output last layers of the densitygrad* / densitycurvature* modes (o[0] density, o[1..3] predicted gradient)
    // ADJ: keep the pre-activation accumulators of every activated layer for the adjoint pass (the reference stores them in
    // sStorageForAdjointHidden, :912-920, 997-1006)
    static constexpr int MAXL = 8;
    struct Pre { uint32_t v[MAXL][2][NT][2]; };
    template <bool ADJ = false>
    static __device__ __forceinline__ void forward(const NetParams& P, uint32_t sblob, uint4* stage, const Act& ac,
                                                   float px, float py, float pz, float3 dir, float tfrac, bool four, __half (&o)[4], Pre* pre = nullptr) {
        const int lane = threadIdx.x & 31;
        __syncwarp();
        uint32_t row[K0 / 2];
        build_row(P, px, py, pz, dir, tfrac, row);
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
            Core::template dense<KT0>(acc, sblob + P.off_w0, bias_addr, lane, [&](int kt, uint32_t (&a)[2][4]) {
#pragma unroll
                for (int mt = 0; mt < 2; ++mt)
                    ldmatrix_x4(a[mt], stage_addr + ((mt * 16 + lrow) * RS_CHUNKS + kt * 2 + lcol) * 16);
            });
        }
        if constexpr (ADJ) copy_frag(pre->v[0], acc);
        activate(acc, P.activation, ac);
#pragma unroll 1
        for (int l = 0; l < P.num_hidden; ++l) {
            uint32_t nxt[2][NT][2];
            Core::template dense<KTH>(nxt, sblob + P.off_wh + l * Core::wh_bytes(), bias_addr + (1 + l) * H * 2, lane,
                                      [&](int kt, uint32_t (&a)[2][4]) {
#pragma unroll
                                          for (int mt = 0; mt < 2; ++mt) {
                                              a[mt][0] = acc[mt][2 * kt][0]; a[mt][1] = acc[mt][2 * kt][1];
                                              a[mt][2] = acc[mt][2 * kt + 1][0]; a[mt][3] = acc[mt][2 * kt + 1][1];
                                          }
                                      });
            if constexpr (ADJ) copy_frag(pre->v[1 + l], nxt);
            activate(nxt, P.activation, ac);
            copy_frag(acc, nxt);
        }
        // activations back to one row per sample (the reference's store_matrix_sync to shared memory, :1018-1023):
        // accumulator register j of tile (mt, nt) holds sample mt*16 + lane/4 + 8*j, channels nt*8 + 2*(lane%4) + {0,1}
        __syncwarp();
        frag_to_rows<NT>(stage, acc, lane);
        __syncwarp();
        if (four) {
            // :1045-1052 (4 outputs) and :1106-1115 (6 outputs, the first four laid out the same): out1 = (bias0, bias1),
            // out2 = (bias2, bias3); per hidden channel v: out1 = hfma2((v,v), (w0,w1), out1), out2 = hfma2((v,v), (w2,w3), out2)
            const uint32_t rows = sblob + P.off_wlast_rows;
            const uint2 bl = lds64(bias_addr + (1 + P.num_hidden) * H * 2);
            __half2 out1 = u2h(bl.x), out2 = u2h(bl.y);
#pragma unroll
            for (int c4 = 0; c4 < H / 8; ++c4) {
                const uint4 v = stage[lane * RS_CHUNKS + c4];
                const uint32_t u[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const __half lo = __low2half(u2h(u[k])), hi = __high2half(u2h(u[k]));
                    const uint2 w0 = lds64(rows + (c4 * 8 + 2 * k) * 16), w1 = lds64(rows + (c4 * 8 + 2 * k + 1) * 16);
                    out1 = __hfma2(__halves2half2(lo, lo), u2h(w0.x), out1); out2 = __hfma2(__halves2half2(lo, lo), u2h(w0.y), out2);
                    out1 = __hfma2(__halves2half2(hi, hi), u2h(w1.x), out1); out2 = __hfma2(__halves2half2(hi, hi), u2h(w1.y), out2);
                }
            }
            o[0] = __low2half(out1); o[1] = __high2half(out1); o[2] = __low2half(out2); o[3] = __high2half(out2);
            return;
        }
        // last layer, :1138-1143: out = bias; out = hfma(x[c], w[c], out) for c = 0..H-1
        const __half* wl = reinterpret_cast<const __half*>(P.wlast2);
        __half out = __float2half(P.blast);
#pragma unroll
        for (int c4 = 0; c4 < H / 8; ++c4) {
            const uint4 v = stage[lane * RS_CHUNKS + c4];
            const uint32_t u[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                out = __hfma(__low2half(u2h(u[k])), wl[c4 * 8 + 2 * k], out);
                out = __hfma(__high2half(u2h(u[k])), wl[c4 * 8 + 2 * k + 1], out);
            }
        }
        o[0] = out; o[1] = o[2] = o[3] = __float2half(0.f);
    }

    // ------------------------------------------------------------------------------------------------------------------
    // GRADIENT_MODE 2: d density / d position by one backward pass (_evalNormalAdjoint, renderer_volume_tensorcores.cuh:
    // 1205-1530), the reference's operations in the reference's order. Samples stay the M dimension of the MMA: the reference
    // computes adj_in(cin x 32 samples) = W^T * adj_out with wmma, this is its transpose - the same dot products, fp16
    // accumulators from zero, K (= cout) consumed 16 at a time in ascending order.
    struct AdjAct { __half p, q2, inv_q2, one; };
    static __device__ __forceinline__ AdjAct make_adj_act(float param) {
        AdjAct a;
        a.p = __float2half(param);
        const float p2 = rf::add(param, param);
        a.q2 = __float2half(p2);                    // Snake / SnakeAlt: half(2 * param), renderer_activations.cuh:299, 368
        a.inv_q2 = __float2half(rf::rcp(p2));       // SnakeAlt: half(1 / (2 * param)), :369 (fast-math reciprocal)
        a.one = __float2half(1.0f);
        return a;
    }
    // ACTIVATION::adjoint(x, zAdj) of the half overloads, the __CUDA_ARCH__ < 800 branches (the reference builds for compute_75)
    static __device__ __forceinline__ __half act_adj1(int act, const AdjAct& a, __half x, __half z) {
        if (act == D_ACT_RELU) return __half2float(x) > 0.f ? z : __float2half(0.f);                         // :66-73
        if (act == D_ACT_SINE) return __hmul(z, __hmul(a.p, exact_hcos(__hmul(x, a.p))));                         // :123-127
        if (act == D_ACT_SNAKE) return __hmul(z, __hadd(a.one, exact_hsin(__hmul(a.q2, x))));                     // :295-300
        if (act == D_ACT_SNAKEALT) return __hmul(z, __hadd(exact_hsin(__hmul(a.q2, x)), a.inv_q2));               // :366-371
        return z;
    }
    static __device__ __forceinline__ void act_adjoint(uint32_t (&z)[2][NT][2], const uint32_t (&x)[2][NT][2], int act, const AdjAct& a) {
#pragma unroll
        for (int mt = 0; mt < 2; ++mt)
#pragma unroll
            for (int nt = 0; nt < NT; ++nt)
#pragma unroll
                for (int j = 0; j < 2; ++j) {
                    const __half2 zv = u2h(z[mt][nt][j]), xv = u2h(x[mt][nt][j]);
                    z[mt][nt][j] = h2u(__halves2half2(act_adj1(act, a, __low2half(xv), __low2half(zv)), act_adj1(act, a, __high2half(xv), __high2half(zv))));
                }
    }
    // out(32 x 8*NTL) = X(32 x H) * B, B as fragments at w_addr (api.cu build_blob, adjoint sections)
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
    // one axis of LoadVolumetricFeaturesAdjoint (:660-678): central difference of the time-interpolated texel, dotted with the adjoint
    static __device__ __forceinline__ float latent_axis(const NetParams& P, int q, float xh, float yh, float zh, float xl, float yl, float zl,
                                                        float tfrac, float s, const float (&adj)[4]) {
        const float4 ah = tex3D<float4>(P.tex_a[q], xh, yh, zh), al = tex3D<float4>(P.tex_a[q], xl, yl, zl);
        const float4 bh = tex3D<float4>(P.tex_b[q], xh, yh, zh), bl = tex3D<float4>(P.tex_b[q], xl, yl, zl);
        const float hx = rf::fma(tfrac, rf::sub(bh.x, ah.x), ah.x), hy = rf::fma(tfrac, rf::sub(bh.y, ah.y), ah.y);
        const float hz = rf::fma(tfrac, rf::sub(bh.z, ah.z), ah.z), hw = rf::fma(tfrac, rf::sub(bh.w, ah.w), ah.w);
        const float lx = rf::fma(tfrac, rf::sub(bl.x, al.x), al.x), ly = rf::fma(tfrac, rf::sub(bl.y, al.y), al.y);
        const float lz = rf::fma(tfrac, rf::sub(bl.z, al.z), al.z), lw = rf::fma(tfrac, rf::sub(bl.w, al.w), al.w);
        const float gx = rf::mul(s, rf::sub(hx, lx)), gy = rf::mul(s, rf::sub(hy, ly)), gz = rf::mul(s, rf::sub(hz, lz)), gw = rf::mul(s, rf::sub(hw, lw));
        return rf::fma(adj[3], gw, rf::fma(adj[2], gz, rf::fma(adj[0], gx, rf::mul(adj[1], gy))));      // dot(adjC, grad) as compiled
    }

    // all 32 lanes call, right after forward<true> of the same position; out_last = the raw density output of this lane's sample
    static __device__ __forceinline__ float3 adjoint(const NetParams& P, uint32_t sblob, uint4* stage, const AdjAct& aa, float px, float py, float pz,
                                                     float tfrac, bool sigm, __half out_last, const Pre& pre) {
        const int lane = threadIdx.x & 31;
        // adjoint of the output parametrisation, :1233-1246: sigmoid -> Sigmoid::adjoint(float) = ev / (ev + 1)^2 in double; else 1
        __half adj_out = __float2half(1.0f);
        if (sigm) {
            const float ev = rf::ex2(rf::mul(__half2float(out_last), 1.4426950408889634f));
            const double e = double(ev), e1 = e + 1.0;
            adj_out = __float2half(float(e / (e1 * e1)));
        }
        // adjoint of the last layer, :1248-1262: row[cin] = hmul(wLast[cin] (density output), adjOut)
        __syncwarp();
        uint32_t* st32 = reinterpret_cast<uint32_t*>(stage);
        const __half2 adj2 = __halves2half2(adj_out, adj_out);
#pragma unroll
        for (int c2 = 0; c2 < H / 2; ++c2) {
            // wlast2 = the weights of output 0 (the density; sWeightsLast[cin] or sWeightsLast[2 * cin].x of the 4-output layers)
            st32[lane * (RS_CHUNKS * 4) + c2] = h2u(__hmul2(u2h(P.wlast2[c2]), adj2));
        }
        __syncwarp();
        uint32_t z[2][NT][2];
        rows_to_frag(stage, z, lane);
        // hidden layers, last to first, :1272-1334
#pragma unroll 1
        for (int l = P.num_hidden - 1; l >= 0; --l) {
            act_adjoint(z, pre.v[1 + l], P.activation, aa);
            uint32_t nxt[2][NT][2];
            dense_t<NT>(nxt, sblob + P.off_adj_wh + l * Core::wh_bytes(), lane, z);
            copy_frag(z, nxt);
        }
        // layer 0, :1340-1448
        act_adjoint(z, pre.v[0], P.activation, aa);
        float3 out = make_float3(0.f, 0.f, 0.f);
        const float h = P.fd_stepsize;                      // latentGridDifferencesStepSize
        const float s = rf::rcp(rf::add(h, h));             // 1 / (2 h)
        const float xh = rf::add(px, h), xl = rf::sub(px, h), yh = rf::add(py, h), yl = rf::sub(py, h), zh = rf::add(pz, h), zl = rf::sub(pz, h);
#pragma unroll
        for (int g = 0; g < GC16; ++g) {
            uint32_t c16[2][2][2];
            dense_t<2>(c16, sblob + P.off_adj_w0g + g * (H * 16 * 2), lane, z);
            __syncwarp();
            frag_to_rows<2>(stage, c16, lane);
            __syncwarp();
            const uint4 r0 = stage[lane * RS_CHUNKS], r1 = stage[lane * RS_CHUNKS + 1];
            const uint32_t u[8] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w};
            float3 adjp = make_float3(0.f, 0.f, 0.f);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const float adj[4] = {__half2float(__low2half(u2h(u[2 * i]))), __half2float(__high2half(u2h(u[2 * i]))),
                                      __half2float(__low2half(u2h(u[2 * i + 1]))), __half2float(__high2half(u2h(u[2 * i + 1])))};
                const int q = 4 * g + i;
                adjp.x = rf::add(adjp.x, latent_axis(P, q, xh, py, pz, xl, py, pz, tfrac, s, adj));
                adjp.y = rf::add(adjp.y, latent_axis(P, q, px, yh, pz, px, yl, pz, tfrac, s, adj));
                adjp.z = rf::add(adjp.z, latent_axis(P, q, px, py, zh, px, py, zl, tfrac, s, adj));
            }
            out.x = rf::add(out.x, adjp.x); out.y = rf::add(out.y, adjp.y); out.z = rf::add(out.z, adjp.z);
        }
        // the first H inputs: position, (time), Fourier features, :1418-1446 and :1466-1504
        {
            uint32_t zf[2][NT][2];
            dense_t<NT>(zf, sblob + P.off_adj_w0f, lane, z);
            __syncwarp();
            frag_to_rows<NT>(stage, zf, lane);
            __syncwarp();
        }
        const __half* row = reinterpret_cast<const __half*>(stage + lane * RS_CHUNKS);
        out.x = rf::add(out.x, __half2float(row[0])); out.y = rf::add(out.y, __half2float(row[1])); out.z = rf::add(out.z, __half2float(row[2]));
        const __half vx = __float2half(px), vy = __float2half(py), vz = __float2half(pz);
        const __half* F = reinterpret_cast<const __half*>(P.fourier2);
#pragma unroll 2
        for (int i = 0; i < NF; ++i) {
            __half c = __hmul(vx, F[i]);
            c = __hfma(vy, F[i + NF], c);
            c = __hfma(vz, F[i + 2 * NF], c);
            // hsub(hmul(adjSin, hcos(c)), hmul(adjCos, hsin(c))) in the source (:1486-1495); cuda_fp16.hpp's mul.f16 / sub.f16 carry no
            // rounding modifier, and the reference's compiled kernel fuses the first product into the subtraction (SASS of the reference
            // PTX: HMUL2 t = adjCos * sin; HFMA2 adjSin * cos - t) - written out here so that it does not depend on this file's inlining
            const __half adj_c = __hfma(row[FOFF + NF + i], exact_hcos(c), __hneg(__hmul_rn(row[FOFF + i], exact_hsin(c))));
            out.x = rf::add(out.x, __half2float(__hmul(F[i], adj_c)));
            out.y = rf::add(out.y, __half2float(__hmul(F[i + NF], adj_c)));
            out.z = rf::add(out.z, __half2float(__hmul(F[i + 2 * NF], adj_c)));
        }
        __syncwarp();
        return out;
    }
};

// transfer functions in the reference's operation order
__device__ __forceinline__ float4 exact_tf_piecewise(const float* tf, int R, float d, float stepsize) {   // renderer_tf_piecewise.cuh:27-62
    int i = 0;
    for (; i < R - 2; ++i)
        if (tf[(i + 1) * 5 + 4] > d) break;
    const float* v0 = tf + i * 5;
    const float* v1 = v0 + 5;
    const float p0 = v0[4], p1 = v1[4];
    const float dc = rf::fmax(p0, rf::fmin(d, p1));
    const float f = rf::div(rf::sub(dc, p0), rf::sub(p1, p0));
    float4 c;
    c.x = rf::fma(rf::sub(v1[0], v0[0]), f, v0[0]);
    c.y = rf::fma(rf::sub(v1[1], v0[1]), f, v0[1]);
    c.z = rf::fma(rf::sub(v1[2], v0[2]), f, v0[2]);
    c.w = rf::mul(stepsize, rf::fma(rf::sub(v1[3], v0[3]), f, v0[3]));
    return c;
}
__device__ __forceinline__ float4 exact_tf_texture(const float* tf, int R, float d, float stepsize) {     // renderer_tf_texture.cuh:46-56
    const float x = rf::fma(d, float(R), -0.5f);
    const float fl = rf::floor(x);
    const int di = int(fl);
    const float df = rf::sub(x, float(di));
    const float* v0 = tf + 4 * min(max(di, 0), R - 1);
    const float* v1 = tf + 4 * min(max(di + 1, 0), R - 1);
    float4 c;
    c.x = rf::fma(df, rf::sub(v1[0], v0[0]), v0[0]);
    c.y = rf::fma(df, rf::sub(v1[1], v0[1]), v0[1]);
    c.z = rf::fma(df, rf::sub(v1[2], v0[2]), v0[2]);
    c.w = rf::mul(rf::fma(df, rf::sub(v1[3], v0[3]), v0[3]), stepsize);
    return c;
}

// one sample of the reference marcher after volume.eval: density range + TF (renderer_ray_evaluation_stepping_dvr.cuh:114-133)
__device__ __forceinline__ float4 exact_shade(const DvrParams& D, const float* tfp, float value, float inv_range, bool valid) {
    float4 c = make_float4(0.f, 0.f, 0.f, 0.f);
    if (valid && value >= D.dmin) {
        const float d2 = rf::fmax(0.f, rf::fmin(rf::mul(inv_range, rf::sub(value, D.dmin)), 1.0f));
        if (D.tf_kind == 2) {       // TransferFunctionIdentity::eval, renderer_tf_identity.cuh:36-55: tfp = {scale absorption, scale emission}
            const float e = rf::mul(d2, tfp[1]);
            c = make_float4(e, e, e, rf::mul(rf::mul(d2, tfp[0]), D.stepsize));
        } else
        c = D.tf_kind == 0 ? exact_tf_piecewise(tfp, D.tf_R, d2, D.stepsize) : exact_tf_texture(tfp, D.tf_R, d2, D.stepsize);
    }
    return c;
}
// Blending::eval, renderer_blending.cuh:36-51
__device__ __forceinline__ void exact_blend(const DvrParams& D, float4 c, float t, bool valid, float4& acc, float& acc_depth) {
    if (c.w > 0.f && valid) {
        const float a = (D.blend_mode == 1) ? rf::sub(1.0f, rf::ex2(rf::mul(c.w, -1.4426950408889634f))) : rf::fmin(1.0f, c.w);
        const float om = rf::sub(1.0f, acc.w);
        acc.x = rf::fma(rf::mul(om, c.x), a, acc.x);
        acc.y = rf::fma(rf::mul(om, c.y), a, acc.y);
        acc.z = rf::fma(rf::mul(om, c.z), a, acc.z);
        acc.w = rf::fma(om, a, acc.w);
        acc_depth = rf::fma(rf::mul(om, t), a, acc_depth);
    }
}
// pixel write, renderer_image_evaluator_simple.cuh:105-124: depth /= alpha first, then colour and normal /= samples
__device__ __forceinline__ void exact_write(const DvrParams& D, int b, int x, int y, float4 pix, float pix_depth, int last_steps,
                                            float3 pix_n = make_float3(0.f, 0.f, 0.f)) {
    const float ns = float(D.spp);
    if (D.step_count) D.step_count[((size_t)b * D.H + y) * D.W + x] = last_steps;
    const float3 nrm = make_float3(rf::div(pix_n.x, ns), rf::div(pix_n.y, ns), rf::div(pix_n.z, ns));
    if (D.out_format != 0) {
        store_rgba(D, b, x, y, select_channel(D.channel_mode, D.tonemap, D.max_exposure,
                                              make_float4(rf::div(pix.x, ns), rf::div(pix.y, ns), rf::div(pix.z, ns), rf::div(pix.w, ns)), nrm));
        return;
    }
    float* o = D.out + (long long)b * D.sb + (long long)y * D.sy + (long long)x * D.sx;
    o[0 * D.sc] = rf::div(pix.x, ns); o[1 * D.sc] = rf::div(pix.y, ns); o[2 * D.sc] = rf::div(pix.z, ns);
    o[3 * D.sc] = rf::div(pix.w, ns);
    o[4 * D.sc] = nrm.x; o[5 * D.sc] = nrm.y; o[6 * D.sc] = nrm.z;
    o[7 * D.sc] = rf::div(pix_depth, pix.w);
}

// ---- normals and shading, operation order of the reference kernels oracle/_ref/h32x4_relu_densgrad*.ptx /
// h32x4_relu_fd_phong.ptx
// Sigmoid::forward under fast math (renderer_activations.cuh:174): rcp(ex2(-x * log2 e) + 1)
__device__ __forceinline__ float exact_sigmoid(float x) { return rf::rcp(rf::add(rf::ex2(rf::mul(x, -1.4426950408889634f)), 1.0f)); }
__device__ __forceinline__ float exact_dot(float3 a, float3 b) { return rf::fma(a.z, b.z, rf::fma(a.x, b.x, rf::mul(a.y, b.y))); }
// safeNormalize (helper_math.cuh:2443-2448; the threshold compare is done in double): also returns |v|^2 and rsqrt(|v|^2)
__device__ __forceinline__ float3 exact_safe_normalize(float3 v, float& len2, float& inv) {
    len2 = exact_dot(v, v);
    inv = rf::rsqrt(len2);
    const bool small = double(len2) < 1e-8;
    return make_float3(small ? v.x : rf::mul(v.x, inv), small ? v.y : rf::mul(v.y, inv), small ? v.z : rf::mul(v.z, inv));
}
// BRDFLambert::eval, renderer_brdf_lambert.cuh:57-103; n = safeNormalize(g), g2 = |g|^2, ginv = rsqrt(g2) ("gradientNorm", :63)
__device__ __forceinline__ float4 exact_brdf(const DvrParams& D, float4 c, float3 wpos, float3 n, float g2, float ginv, float3 dir) {
    float absorption = c.w;
    if (D.brdf_flags & 1) absorption = rf::mul(c.w, rf::sub(1.0f, rf::ex2(rf::mul(rf::mul(g2, D.brdf_mag_scaling), -1.4426950408889634f))));
    if (D.brdf_flags & 2) {
        float3 l;           // lightDirection
        float ndl, dnl;     // dot(normal, l) and dot(-normal, l)
        if (D.light_type == 1) {    // directional: normalize(-lightParameter)
            const float3 L = make_float3(D.light[0], D.light[1], D.light[2]);
            const float linv = rf::rsqrt(exact_dot(L, L));
            const float px = rf::mul(L.x, linv), py = rf::mul(L.y, linv), pz = rf::mul(L.z, linv);
            l = make_float3(rf::neg(px), rf::neg(py), rf::neg(pz));
            const float ny = rf::mul(n.y, py);
            ndl = rf::sub(rf::sub(rf::mul(n.x, l.x), ny), rf::mul(n.z, pz));
            dnl = rf::fma(n.z, pz, rf::fma(n.x, px, ny));
        } else {                    // point: normalize(lightParameter - position)
            const float3 d = make_float3(rf::sub(D.light[0], wpos.x), rf::sub(D.light[1], wpos.y), rf::sub(D.light[2], wpos.z));
            const float linv = rf::rsqrt(exact_dot(d, d));
            l = make_float3(rf::mul(d.x, linv), rf::mul(d.y, linv), rf::mul(d.z, linv));
            const float a = rf::mul(n.x, l.x), b = rf::mul(n.y, l.y), e = rf::mul(n.z, l.z);
            ndl = rf::add(e, rf::add(a, b));
            dnl = rf::sub(rf::sub(rf::neg(b), a), e);
        }
        const float lo = rf::sub(D.brdf_mag_center, D.brdf_mag_radius);
        const float range = rf::sub(rf::add(D.brdf_mag_center, D.brdf_mag_radius), lo);
        const float y = rf::fmax(0.f, rf::fmin(rf::div(rf::sub(ginv, lo), range), 1.0f));
        const float phong = rf::mul(rf::mul(y, y), rf::sub(3.0f, rf::add(y, y)));                // smoothstep
        const float amb = rf::fma(rf::add(D.brdf_ambient, -1.0f), phong, 1.0f);                 // lerp(1, ambient, phong)
        const float3 r = make_float3(rf::sub(l.x, rf::mul(rf::mul(n.x, -2.0f), dnl)), rf::sub(l.y, rf::mul(rf::mul(n.y, -2.0f), dnl)),
                                     rf::sub(l.z, rf::mul(rf::mul(n.z, -2.0f), dnl)));          // reflect(l, -normal)
        float rd = exact_dot(dir, r);
        rd = (rd < 0.f) ? 0.f : rd;
        const float e = float(D.brdf_spec_exp);
        const float coef = float(double(rf::add(e, 2.0f)) * 0.159155);
        const float spec = rf::mul(D.brdf_specular, rf::mul(rf::ex2(rf::mul(rf::lg2(rd), e)), coef));
        const float om = rf::sub(1.0f, amb), d = rf::abs(ndl);
        c.x = rf::fma(c.x, amb, rf::mul(om, rf::fma(c.x, d, spec)));
        c.y = rf::fma(c.y, amb, rf::mul(om, rf::fma(c.y, d, spec)));
        c.z = rf::fma(c.z, amb, rf::mul(om, rf::fma(c.z, d, spec)));
    }
    c.w = absorption;
    return c;
}
// Blending::eval with the normal channel, renderer_blending.cuh:36-51
__device__ __forceinline__ void exact_blend_normal(const DvrParams& D, float4 c, float3 n, float t, float4& acc, float3& acc_n, float& acc_depth) {
    const float a = (D.blend_mode == 1) ? rf::sub(1.0f, rf::ex2(rf::mul(c.w, -1.4426950408889634f))) : rf::fmin(1.0f, c.w);
    const float om = rf::sub(1.0f, acc.w);
    acc.x = rf::fma(rf::mul(om, c.x), a, acc.x);
    acc.y = rf::fma(rf::mul(om, c.y), a, acc.y);
    acc.z = rf::fma(rf::mul(om, c.z), a, acc.z);
    acc_n.x = rf::fma(rf::mul(om, n.x), a, acc_n.x);
    acc_n.y = rf::fma(rf::mul(om, n.y), a, acc_n.y);
    acc_n.z = rf::fma(rf::mul(om, n.z), a, acc_n.z);
    acc_depth = rf::fma(rf::mul(om, t), a, acc_depth);
    acc.w = rf::fma(om, a, acc.w);
}

// ADJ: the instantiation that can run the adjoint pass (gradient_mode 2); it keeps the pre-activations of every forward pass
template <int H16, int GC16 = 1, bool DIR = false, bool ADJ = false>
__global__ void __launch_bounds__(kWarpsPerBlock * 32)
dvr_exact_kernel(const __grid_constant__ NetParams P, const __grid_constant__ DvrParams D) {
    using Net = SrnExact<H16, GC16, DIR>;
    using Core = typename Net::Core;
    extern __shared__ uint4 smem[];
    uint4* sblob = smem;
    const int blob16 = (P.blob_bytes + 15) / 16;
    float* stf = reinterpret_cast<float*>(smem + blob16);
    uint4* stage_all = smem + blob16 + kTfSmemFloats / 4;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint4* stage = stage_all + warp * (Core::STAGE_BYTES / 16);

    load_blob(P, sblob);
    const int tf_floats = D.tf_kind == 0 ? D.tf_R * 5 : (D.tf_kind == 1 ? D.tf_R * 4 : 2);
    if (D.tf_batch == 1)
        for (int i = threadIdx.x; i < tf_floats; i += blockDim.x) stf[i] = __ldg(D.tf + i);
    __syncthreads();

    const uint32_t sblob_addr = smem_u32(sblob);
    const typename Net::Act ac = Net::make_act(P.activation, P.act_param);
    const typename Net::AdjAct aa = Net::make_adj_act(P.act_param);
    typename Net::Pre pre;
    const float3 bmin = make_float3(P.box_min[0], P.box_min[1], P.box_min[2]);
    const float3 bsize = make_float3(P.box_size[0], P.box_size[1], P.box_size[2]);
    const float tfrac = rf::sub(P.time_value, rf::floor(P.time_value));       // fracf(time), :596-597
    const float inv_range = rf::rcp(rf::sub(D.dmax, D.dmin));                 // 1 / (densityMax - densityMin), fast math
    const bool four = P.output_mode >= D_OUT_DENSITY_GRADIENT;                // 4 / 6-output last layer
    const bool sigm = output_mode_has_sigmoid(P.output_mode);
    unsigned long long n_valid = 0, n_exec = 0;
    for (;;) {
        int tile = 0;
        if (lane == 0) tile = int(atomicAdd(D.tile_counter, 1u));
        tile = __shfl_sync(0xffffffffu, tile, 0);
        if (tile >= D.num_tiles) break;
        const int tiles_per_img = D.tiles_x * D.tiles_y;
        const int bl = tile / tiles_per_img;
        const int trem = tile - bl * tiles_per_img;
        const int ty = trem / D.tiles_x, tx = trem - ty * D.tiles_x;
        const int b = D.b0 + bl;
        const int x = D.x0 + tx * kTileW + (lane & (kTileW - 1));
        const int y = D.y0 + ty * kTileH + (lane / kTileW);
        const bool in_img = (x < D.x0 + D.tw) && (y < D.y0 + D.th);
        const float* tfp = (D.tf_batch == 1) ? stf : (D.tf + (size_t)b * tf_floats);
        const CamFrame cf = load_camera(D, b);

        float4 pix = make_float4(0.f, 0.f, 0.f, 0.f);
        float3 pix_n = make_float3(0.f, 0.f, 0.f);
        float pix_depth = 0.f;
        int last_steps = 0;
        for (int s = 0; s < D.spp; ++s) {
            const Ray r = make_ray(D, cf, x, y, s, bmin, bsize, D.spp);
            const float tmax = in_img ? r.tmax : -1.f;
            float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
            float3 acc_n = make_float3(0.f, 0.f, 0.f);
            float acc_depth = 0.f;
            int steps = 0;
            for (int i = 0;; ++i) {     // renderer_ray_evaluation_stepping_dvr.cuh:84-154
                const float t = rf::fma(D.stepsize, float(i), r.tmin);
                bool valid = (t <= tmax);
                if (D.early_out) valid = valid && (acc.w < D.alpha_early_out);
                if (!__any_sync(0xffffffffu, valid)) break;
                n_exec += 32;
                const float wx = rf::fma(r.dir.x, t, r.eye.x), wy = rf::fma(r.dir.y, t, r.eye.y), wz = rf::fma(r.dir.z, t, r.eye.z);
                const float px = rf::div(rf::sub(wx, bmin.x), bsize.x), py = rf::div(rf::sub(wy, bmin.y), bsize.y),
                            pz = rf::div(rf::sub(wz, bmin.z), bsize.z);
                __half o[4];
                Net::template forward<ADJ>(P, sblob_addr, stage, ac, px, py, pz, r.dir, tfrac, four, o, &pre);
                float value = __half2float(o[0]);
                if (sigm) value = exact_sigmoid(value);
                steps += valid ? 1 : 0;
                if (!D.normals) {
                    const float4 c = exact_shade(D, tfp, value, inv_range, valid);
                    exact_blend(D, c, t, valid, acc, acc_depth);
                    continue;
                }
                // normals, :114-153: evaluated when ANY lane of the warp shades its sample
                const bool any_shaded = __any_sync(0xffffffffu, valid && value >= D.dmin);
                float3 g = make_float3(0.f, 0.f, 0.f);
                if (P.gradient_mode == 1) {     // central differences of valueNoClamping, renderer_volume_tensorcores.cuh:1185-1197
                    if (any_shaded) {
                        const float h = P.fd_stepsize;
                        float d[6];
#pragma unroll 1
                        for (int k = 0; k < 6; ++k) {
                            // position +- h e_axis: the "+" samples add 0 to the other two coordinates, the "-" samples leave them alone
                            const int axis = k >> 1;
                            const bool plus = (k & 1) == 0;
                            const float qx = axis == 0 ? (plus ? rf::add(wx, h) : rf::sub(wx, h)) : (plus ? rf::add(wx, 0.f) : wx);
                            const float qy = axis == 1 ? (plus ? rf::add(wy, h) : rf::sub(wy, h)) : (plus ? rf::add(wy, 0.f) : wy);
                            const float qz = axis == 2 ? (plus ? rf::add(wz, h) : rf::sub(wz, h)) : (plus ? rf::add(wz, 0.f) : wz);
                            __half q[4];
                            Net::forward(P, sblob_addr, stage, ac, rf::div(rf::sub(qx, bmin.x), bsize.x), rf::div(rf::sub(qy, bmin.y), bsize.y),
                                         rf::div(rf::sub(qz, bmin.z), bsize.z), r.dir, tfrac, four, q);
                            float v = __half2float(q[0]);
                            if (sigm) v = exact_sigmoid(v);
                            if (k == 0) d[0] = v; else if (k == 1) d[1] = v; else if (k == 2) d[2] = v;
                            else if (k == 3) d[3] = v; else if (k == 4) d[4] = v; else d[5] = v;
                        }
                        const float h2 = rf::add(h, h);
                        g = make_float3(rf::div(rf::sub(d[0], d[1]), h2), rf::div(rf::sub(d[2], d[3]), h2), rf::div(rf::sub(d[4], d[5]), h2));
                    }
                } else if (ADJ && P.gradient_mode == 2) {     // the backward pass through the network, :1199-1201
                    if constexpr (ADJ) {
                        // (measured: repeating the forward pass on shaded steps only, instead of keeping the pre-activations of
                        // every step, is slower - 133.6 vs 112.8 ms per C3 frame: nearly every step of a warp shades something)
                        if (any_shaded) g = Net::adjoint(P, sblob_addr, stage, aa, px, py, pz, tfrac, sigm, o[0], pre);
                    }
                } else if (four && any_shaded) {
                    g = make_float3(__half2float(o[1]), __half2float(o[2]), __half2float(o[3]));
                    if (P.output_mode == D_OUT_DENSITY_GRADIENT_CUBIC)
                        g = make_float3(rf::mul(rf::mul(g.x, g.x), g.x), rf::mul(rf::mul(g.y, g.y), g.y), rf::mul(rf::mul(g.z, g.z), g.z));
                }
                float4 c = exact_shade(D, tfp, value, inv_range, valid);
                if (c.w > 0.f) {
                    float g2, ginv;
                    const float3 n = exact_safe_normalize(g, g2, ginv);
                    if (D.brdf_flags) c = exact_brdf(D, c, make_float3(wx, wy, wz), n, g2, ginv, r.dir);
                    if (valid) exact_blend_normal(D, c, n, t, acc, acc_n, acc_depth);
                }
            }
            n_valid += steps;
            last_steps = steps;
            pix.x = rf::add(pix.x, acc.x); pix.y = rf::add(pix.y, acc.y); pix.z = rf::add(pix.z, acc.z); pix.w = rf::add(pix.w, acc.w);
            pix_n.x = rf::fma(acc.w, acc_n.x, pix_n.x); pix_n.y = rf::fma(acc.w, acc_n.y, pix_n.y); pix_n.z = rf::fma(acc.w, acc_n.z, pix_n.z);
            pix_depth = rf::fma(acc.w, acc_depth, pix_depth);       // renderer_image_evaluator_simple.cuh:101-103
        }
        if (in_img) exact_write(D, b, x, y, pix, pix_depth, last_steps, pix_n);
    }
    if (D.counters) {
        for (int off = 16; off > 0; off >>= 1) n_valid += __shfl_xor_sync(0xffffffffu, n_valid, off);
        if (lane == 0) { atomicAdd(D.counters + 0, n_valid); atomicAdd(D.counters + 1, n_exec); }
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// Ground-truth density grid through the same marcher: VolumeInterpolationGrid::eval, texture path, trilinear
// (renderer_volume_grid.cuh:110-139, 196-231) - what the reference renders its comparison images with. The reference traces
// these rays per thread (KERNEL_SYNCHRONIZED_TRACING 0); operation order as in its compiled kernel.
#ifndef FVSRN_EXACT_WITHOUT_GRID_KERNELS     // (the second translation unit of this header, inst_exact_adj.cu, leaves the non-template kernels out)
__device__ __forceinline__ float grid_eval(const GridVolumeParams& G, float wx, float wy, float wz) {
    const float px = rf::mul(rf::div(rf::sub(wx, G.box_min[0]), G.box_size[0]), float(G.rx));
    const float py = rf::mul(rf::div(rf::sub(wy, G.box_min[1]), G.box_size[1]), float(G.ry));
    const float pz = rf::mul(rf::div(rf::sub(wz, G.box_min[2]), G.box_size[2]), float(G.rz));
    return tex3D<float>(G.tex, px, py, pz);
}

__global__ void __launch_bounds__(256)
dvr_grid_kernel(const __grid_constant__ GridVolumeParams G, const __grid_constant__ DvrParams D) {
    __shared__ float stf[kTfSmemFloats];
    const int tf_floats = D.tf_kind == 0 ? D.tf_R * 5 : (D.tf_kind == 1 ? D.tf_R * 4 : 2);
    if (D.tf_batch == 1)
        for (int i = threadIdx.x; i < tf_floats; i += blockDim.x) stf[i] = __ldg(D.tf + i);
    __syncthreads();
    const float3 bmin = make_float3(G.box_min[0], G.box_min[1], G.box_min[2]);
    const float3 bsize = make_float3(G.box_size[0], G.box_size[1], G.box_size[2]);
    const float inv_range = rf::rcp(rf::sub(D.dmax, D.dmin));
    unsigned long long n_valid = 0;
    const long long per_img = (long long)D.tw * D.th, total = per_img * D.nb;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int bl = int(i / per_img);
        const int rem = int(i - (long long)bl * per_img);
        const int y = D.y0 + rem / D.tw, x = D.x0 + rem % D.tw, b = D.b0 + bl;
        const float* tfp = (D.tf_batch == 1) ? stf : (D.tf + (size_t)b * tf_floats);
        const CamFrame cf = load_camera(D, b);
        float4 pix = make_float4(0.f, 0.f, 0.f, 0.f);
        float pix_depth = 0.f;
        int last_steps = 0;
        for (int s = 0; s < D.spp; ++s) {
            const Ray r = make_ray(D, cf, x, y, s, bmin, bsize, D.spp);
            float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
            float acc_depth = 0.f;
            int steps = 0;
            for (int k = 0;; ++k) {
                const float t = rf::fma(D.stepsize, float(k), r.tmin);
                bool valid = (t <= r.tmax);
                if (D.early_out) valid = valid && (acc.w < D.alpha_early_out);
                if (!valid) break;
                const float value = grid_eval(G, rf::fma(r.dir.x, t, r.eye.x), rf::fma(r.dir.y, t, r.eye.y), rf::fma(r.dir.z, t, r.eye.z));
                const float4 c = exact_shade(D, tfp, value, inv_range, true);
                ++steps;
                exact_blend(D, c, t, true, acc, acc_depth);
            }
            n_valid += steps; last_steps = steps;
            pix.x = rf::add(pix.x, acc.x); pix.y = rf::add(pix.y, acc.y); pix.z = rf::add(pix.z, acc.z); pix.w = rf::add(pix.w, acc.w);
            pix_depth = rf::fma(acc.w, acc_depth, pix_depth);
        }
        exact_write(D, b, x, y, pix, pix_depth, last_steps);
    }
    if (D.counters) {
        for (int off = 16; off > 0; off >>= 1) n_valid += __shfl_xor_sync(0xffffffffu, n_valid, off);
        if ((threadIdx.x & 31) == 0) { atomicAdd(D.counters + 0, n_valid); atomicAdd(D.counters + 1, n_valid); }
    }
}

// IVolumeInterpolation::evaluate for the grid: positions (N,3) in the unit cube -> densities (N) (volume_interpolation.cpp:26-127)
__global__ void eval_grid_kernel(const __grid_constant__ GridVolumeParams G, const float* __restrict__ pos, float* __restrict__ out, long long n) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        out[i] = grid_eval(G, pos[3 * i], pos[3 * i + 1], pos[3 * i + 2]);
}
#endif

} // namespace fvsrn
