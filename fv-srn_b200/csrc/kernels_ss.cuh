// Deeper-pipelined variant of the tcgen05 DVR kernel (see kernels_tc.cuh for the protocol): the activations (A operand)
// go through SHARED memory (SS-mode tcgen05.mma: both operands by shared-memory descriptor) instead of TMEM, so a slot
// costs only its fp32 accumulator in TMEM (H columns instead of H + K0/2). With the 512 TMEM columns of an SM split over
// 4 resident CTAs that allows NS = 3 march steps in flight per CTA for hidden width 32 (12 instead of 8 slots per SM):
// a compute warp that publishes one slot has two other slots' epilogues to work on while that slot's MMA chain runs.
//
// A tile of a slot, K-major no-swizzle UMMA layout: K chunk kc (8 halves = 16 B per row) at kc * 2048 B, row r at
// (r/8)*128 + (r%8)*16 = r*16 inside the chunk -> thread t (= row t) writes one STS.128 per chunk at
// a_base + kc*2048 + t*16: a warp writes 512 contiguous bytes, conflict-free. Descriptor: LBO (K direction) = 2048 B,
// SBO (M direction) = 128 B, K=16 step kt starts at a_base + kt*4096.
#pragma once
#include "kernels_tc.cuh"

namespace fvsrn {

template <int NS>
struct SsShared {            // fixed part of the dynamic shared memory, placed first
    unsigned long long ready[NS];  // slot operands written: one arrival per compute warp
    unsigned long long done[NS];   // slot MMA chain complete (tcgen05.commit)
    uint32_t tmem_addr;
    int tile;
    int pad[2];
};

template <int H16, int GC16, int VARIANT, int NS>
struct SrnSs {
    using Tc = SrnTc<H16, GC16, false, VARIANT>;     // row builder, weight-tile layout
    using Core = typename Tc::Core;
    static constexpr int H = Tc::H, K0 = Tc::K0, KT0 = K0 / 16, KTH = H / 16;
    static constexpr bool RELU = Tc::RELU;
    static constexpr int TMEM_USED = NS * H;
    static constexpr int TMEM_COLS = TMEM_USED <= 32 ? 32 : TMEM_USED <= 64 ? 64 : TMEM_USED <= 128 ? 128 : TMEM_USED <= 256 ? 256 : 512;
    static constexpr int A_BYTES = 128 * K0 * 2;     // one slot's A tile (layer 0 uses all K0/8 chunks, hidden layers H/8)
    static constexpr int ONES_BYTES = 128 * 16 * 2;  // constant A tile of the bias step: k = 0 column is 1.0

    struct Ctx {
        uint32_t tmem_lane, tmem_base;
        uint32_t sblob;           // weights
        uint32_t a_tiles, ones;   // shared addresses
        uint32_t ready, done;
        uint32_t phase;
        tc::ActF af;
    };

    static __device__ __forceinline__ uint32_t desc_lo(uint32_t addr, uint32_t lbo) {
        return ((addr >> 4) & 0x3FFFu) | (((lbo >> 4) & 0x3FFFu) << 16);
    }

    // MMA warp (converged): one layer of one slot. A: smem tile of the slot, B: weight tile; then the bias step.
    static __device__ __forceinline__ void issue_layer(const Ctx& c, int slot, uint32_t w_addr, int n, int kt_data, bool bias_chunk) {
        const uint32_t idesc = tc::make_idesc(128, n);
        const uint32_t kstride = uint32_t(n / 8) * 128u;
        const uint32_t hi = (128u >> 4) | (1u << 14);                       // SBO = 128 B, descriptor version 1
        uint32_t b_lo = desc_lo(w_addr, kstride);
        uint32_t a_lo = desc_lo(c.a_tiles + slot * A_BYTES, 2048u);
        const uint32_t d = c.tmem_base + slot * H;
        for (int kt = 0; kt < kt_data; ++kt, b_lo += (2u * kstride) >> 4, a_lo += 4096u >> 4)
            tc::mma_ss2_elect(d, a_lo, hi, b_lo, hi, idesc, kt > 0);
        if (bias_chunk) tc::mma_ss2_elect(d, desc_lo(c.ones, 2048u), hi, b_lo, hi, idesc, 1u);
        tc::mma_commit_elect(c.done + 8 * slot);
    }

    // compute warp: this warp's rows of the slot's A tile are written -> one arrival on the slot's ready barrier
    static __device__ __forceinline__ void publish(const Ctx& c, int slot) {
        asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");      // generic-proxy STS -> async-proxy (MMA) reads
        tc::fence_before_sync();
        __syncwarp();
        if ((threadIdx.x & 31) == 0) tc::mbar_arrive(c.ready + 8 * slot);
    }

    template <int NCHUNK, int N>
    static __device__ __forceinline__ void store_chunks(const Ctx& c, int slot, const uint32_t (&v)[N]) {
        const uint32_t base = c.a_tiles + slot * A_BYTES + threadIdx.x * 16;
#pragma unroll
        for (int kc = 0; kc < NCHUNK; ++kc)
            asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};\n" :: "r"(base + kc * 2048), "r"(v[4 * kc]), "r"(v[4 * kc + 1]),
                         "r"(v[4 * kc + 2]), "r"(v[4 * kc + 3]) : "memory");
    }

    static __device__ __forceinline__ void store_row(const Ctx& c, int slot, const uint32_t (&row)[K0 / 2]) {
        store_chunks<K0 / 8>(c, slot, row);
        publish(c, slot);
    }

    // accumulator row -> activation -> fp16 A tile of the next layer (H == 32 * m)
    static __device__ __forceinline__ void epilogue(const Ctx& c, int slot, int act) {
        static_assert(H % 32 == 0, "SS variant: hidden width must be a multiple of 32");
#pragma unroll
        for (int c0 = 0; c0 < H; c0 += 32) {
            uint32_t v[32];
            tc::tmem_ld32(c.tmem_lane + slot * H + c0, v);
            tc::wait_ld();
            uint32_t pk[16];
#pragma unroll
            for (int j = 0; j < 16; ++j)
                pk[j] = RELU ? tc::pack_relu(__uint_as_float(v[2 * j]), __uint_as_float(v[2 * j + 1])) : tc::act_pack(act, c.af, v[2 * j], v[2 * j + 1]);
            const uint32_t base = c.a_tiles + slot * A_BYTES + (c0 / 8) * 2048 + threadIdx.x * 16;
#pragma unroll
            for (int kc = 0; kc < 4; ++kc)
                asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};\n" :: "r"(base + kc * 2048), "r"(pk[4 * kc]), "r"(pk[4 * kc + 1]),
                             "r"(pk[4 * kc + 2]), "r"(pk[4 * kc + 3]) : "memory");
        }
    }

    // last hidden layer's epilogue of a density network: in-register dot product (see SrnTc::epilogue_dot)
    static __device__ __forceinline__ float epilogue_dot(const NetParams& P, const Ctx& c, int slot, int act) {
        __half2 a0 = __float2half2_rn(0.f), a1 = a0;
#pragma unroll
        for (int c0 = 0; c0 < H; c0 += 32) {
            uint32_t v[32];
            tc::tmem_ld32(c.tmem_lane + slot * H + c0, v);
            tc::wait_ld();
#pragma unroll
            for (int j = 0; j < 16; j += 2) {
                const uint32_t p0 = RELU ? tc::pack_relu(__uint_as_float(v[2 * j]), __uint_as_float(v[2 * j + 1])) : tc::act_pack(act, c.af, v[2 * j], v[2 * j + 1]);
                const uint32_t p1 = RELU ? tc::pack_relu(__uint_as_float(v[2 * j + 2]), __uint_as_float(v[2 * j + 3])) : tc::act_pack(act, c.af, v[2 * j + 2], v[2 * j + 3]);
                a0 = __hfma2(u2h(p0), u2h(P.wlast2[c0 / 2 + j]), a0);
                a1 = __hfma2(u2h(p1), u2h(P.wlast2[c0 / 2 + j + 1]), a1);
            }
        }
        const float2 f0 = __half22float2(a0), f1 = __half22float2(a1);
        return P.blast + ((f0.x + f0.y) + (f1.x + f1.y));
    }

    // compute warps: layer chain of the NS slots (density networks only: the last layer is the in-register dot product)
    static __device__ __forceinline__ void mlp_compute(const NetParams& P, Ctx& c, bool warp_active, float (&out)[NS]) {
        const int act = P.activation;
#pragma unroll 1
        for (int l = 0; l <= P.num_hidden; ++l) {
            const bool dot = l == P.num_hidden;
#pragma unroll
            for (int slot = 0; slot < NS; ++slot) {
                tc::mbar_wait(c.done + 8 * slot, c.phase);
                tc::fence_after_sync();
                if (dot) out[slot] = warp_active ? epilogue_dot(P, c, slot, act) : 0.f;
                else {
                    if (warp_active) epilogue(c, slot, act);
                    publish(c, slot);
                }
            }
            c.phase ^= 1u;
        }
        tc::fence_before_sync();
    }

    // MMA warp: the matching issue schedule of one trip (`c.phase` = parity of the ready barriers)
    static __device__ __forceinline__ void mlp_issue(const NetParams& P, Ctx& c) {
#pragma unroll 1
        for (int l = -1; l < P.num_hidden; ++l) {
#pragma unroll
            for (int slot = 0; slot < NS; ++slot) {
                tc::mbar_wait(c.ready + 8 * slot, c.phase);
                tc::fence_after_sync();
                if (l < 0) issue_layer(c, slot, c.sblob + P.off_w0, H, KT0, P.has_time != 0);
                else issue_layer(c, slot, c.sblob + P.off_wh + l * Tc::wh_bytes(), H, KTH, true);
            }
            c.phase ^= 1u;
        }
    }
};

template <int H16, int GC16, int VARIANT, int NS>
__global__ void __launch_bounds__(kTcThreads, 4)
dvr_ss_kernel(const __grid_constant__ NetParams P, const __grid_constant__ DvrParams D) {
    using Net = SrnSs<H16, GC16, VARIANT, NS>;
    using Tc = typename Net::Tc;
    using Core = typename Net::Core;
    extern __shared__ uint4 smem[];
    SsShared<NS>* sh = reinterpret_cast<SsShared<NS>*>(smem);
    static_assert(sizeof(SsShared<NS>) % 16 == 0, "SsShared must keep the blob 16-byte aligned");
    uint4* sblob = smem + sizeof(SsShared<NS>) / 16;
    const int blob16 = (P.blob_bytes + 15) / 16;
    float* stf = reinterpret_cast<float*>(sblob + blob16);
    uint4* ones = sblob + blob16 + D.tf_smem_floats / 4;
    uint4* a_tiles = ones + Net::ONES_BYTES / 16;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const bool is_mma_warp = warp == 4;

    // ---- prologue: weights, constant ones tile, barriers, TMEM
    load_blob(P, sblob);
    for (int i = threadIdx.x; i < Net::ONES_BYTES / 16; i += blockDim.x)
        ones[i] = (i < 128) ? make_uint4(0x00003C00u, 0u, 0u, 0u) : make_uint4(0u, 0u, 0u, 0u);   // chunk 0: k=0 is 1.0; chunk 1: zeros
    if (threadIdx.x == 0)
        for (int s = 0; s < NS; ++s) { tc::mbar_init(smem_u32(&sh->ready[s]), 4); tc::mbar_init(smem_u32(&sh->done[s]), 1); }
    if (is_mma_warp) tc::tmem_alloc(smem_u32(&sh->tmem_addr), Net::TMEM_COLS);
    if (D.tf_batch == 1) tc_stage_tf(D, 0, stf);
    asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    typename Net::Ctx ctx;
    ctx.tmem_base = sh->tmem_addr;
    ctx.tmem_lane = ctx.tmem_base + (uint32_t((warp & 3) * 32) << 16);
    ctx.sblob = smem_u32(sblob);
    ctx.ones = smem_u32(ones);
    ctx.a_tiles = smem_u32(a_tiles);
    ctx.ready = smem_u32(&sh->ready[0]);
    ctx.done = smem_u32(&sh->done[0]);
    ctx.phase = 0;
    ctx.af = tc::make_actf(P.activation, P.act_param);

    const float3 bmin = make_float3(P.box_min[0], P.box_min[1], P.box_min[2]);
    const float3 bsize = make_float3(P.box_size[0], P.box_size[1], P.box_size[2]);
    const float3 binv = make_float3(__fdiv_rn(1.f, bsize.x), __fdiv_rn(1.f, bsize.y), __fdiv_rn(1.f, bsize.z));
    const int lx = (warp & 1) * 8 + (lane & 7), ly = ((warp >> 1) & 1) * 4 + (lane >> 3);

    unsigned long long n_valid = 0, n_exec = 0;
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
        if (D.tf_batch != 1) {
            tc_stage_tf(D, b, stf);
            __syncthreads();
        }
        if (is_mma_warp) {
            for (int s = 0; s < D.spp; ++s)
                for (;;) {
                    if (!__syncthreads_or(0)) break;
                    Net::mlp_issue(P, ctx);
                }
            continue;
        }
        const int x = D.x0 + tx * kTcTileW + lx;
        const int y = D.y0 + ty * kTcTileH + ly;
        const bool in_img = (x < D.x0 + D.tw) && (y < D.y0 + D.th);
        const CamFrame cf = load_camera(D, b);

        float4 pix_color = make_float4(0.f, 0.f, 0.f, 0.f);
        float pix_depth = 0.f;
        int last_steps = 0;
        for (int s = 0; s < D.spp; ++s) {
            const Ray r = make_ray(D, cf, x, y, s, bmin, bsize, D.spp);
            const float tmax = in_img ? r.tmax : -1.f;
            float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
            float acc_depth = 0.f;
            int steps = 0;
            // NS consecutive steps per trip; steps after the first are speculative, compositing is in order
            for (int i = 0;; i += NS) {
                const bool alive = !D.early_out || (acc.w < D.alpha_early_out);
                float t[NS]; bool in[NS];
#pragma unroll
                for (int k = 0; k < NS; ++k) {
                    t[k] = __fmaf_rn(float(i + k), D.stepsize, r.tmin);
                    in[k] = (t[k] <= tmax) && alive;
                }
                if (!__syncthreads_or(in[0] ? 1 : 0)) break;
                const bool warp_active = __any_sync(0xffffffffu, in[0]);
#pragma unroll
                for (int k = 0; k < NS; ++k) {
                    if (warp_active) {
                        float px = 0.f, py = 0.f, pz = 0.f;
                        if (in[k]) {
                            px = __fmul_rn(__fsub_rn(__fmaf_rn(r.dir.x, t[k], r.eye.x), bmin.x), binv.x);
                            py = __fmul_rn(__fsub_rn(__fmaf_rn(r.dir.y, t[k], r.eye.y), bmin.y), binv.y);
                            pz = __fmul_rn(__fsub_rn(__fmaf_rn(r.dir.z, t[k], r.eye.z), bmin.z), binv.z);
                        }
                        uint32_t row[Core::K0 / 2];
                        Tc::fetch_latent(P, px, py, pz, row);
                        Tc::build_row(P, px, py, pz, r.dir.x, r.dir.y, r.dir.z, row);
                        Net::store_row(ctx, k, row);
                    } else Net::publish(ctx, k);
                }
                float o[NS];
                Net::mlp_compute(P, ctx, warp_active, o);
                if (lane == 0) n_exec += 32 * NS;
                float4 c[NS];
#pragma unroll
                for (int k = 0; k < NS; ++k) { const float ok[4] = {o[k], 0.f, 0.f, 0.f}; c[k] = tc_shade_sample(P, D, stf, ok); }
#pragma unroll
                for (int k = 0; k < NS; ++k) {
                    const bool valid = in[k] && (k == 0 || !D.early_out || (acc.w < D.alpha_early_out));
                    if (valid) {
                        ++steps;
                        blend_sample(D, c[k], t[k], acc, acc_depth);
                    }
                }
            }
            n_valid += steps;
            last_steps = steps;
            pix_color.x += acc.x; pix_color.y += acc.y; pix_color.z += acc.z; pix_color.w += acc.w;
            pix_depth += acc_depth * acc.w;
        }
        if (in_img) write_pixel(D, b, x, y, pix_color, pix_depth, last_steps, D.spp);
    }
    if (D.counters && !is_mma_warp) {
        for (int off = 16; off > 0; off >>= 1) n_valid += __shfl_xor_sync(0xffffffffu, n_valid, off);
        if (lane == 0) { atomicAdd(D.counters + 0, n_valid); atomicAdd(D.counters + 1, n_exec); }
    }
    tc::fence_before_sync();
    __syncthreads();
    if (is_mma_warp) tc::tmem_dealloc(ctx.tmem_base, Net::TMEM_COLS);
}

} // namespace fvsrn
