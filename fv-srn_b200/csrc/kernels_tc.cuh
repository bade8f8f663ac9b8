// tcgen05 / TMEM versions of the two kernels of the hot path (fast mode, see srn_tc.cuh):
//   dvr_tc_kernel   <- ImageEvaluatorSimpleKernel + RayEvaluationSteppingDvr::eval + TF + Blending
//   eval_tc_kernel  <- EvaluateNoBatches
// One CTA = 128 threads = 128 rays (a 16x8 pixel tile, each warp an 8x4 sub-tile) marched in lock-step: one SRN sample
// per thread per step = one M=128 UMMA tile per step. Persistent CTAs pull tiles from an atomic queue; several CTAs
// are resident per SM so that one CTA's MMA / TMEM latency is covered by the others' feature and fetch work.
#pragma once
#include "kernels.cuh"
#include "srn_tc.cuh"

namespace fvsrn {

constexpr int kTcThreads = 128;
constexpr int kTcTileW = 16, kTcTileH = 8;

template <int H16, int GC16, bool DIR>
struct SrnTc {
    using Core = Srn<H16, GC16, DIR>;
    static constexpr int H = Core::H, K0 = Core::K0, KT0 = K0 / 16, KTH = H / 16;
    static constexpr int NLAST = 16;                       // last layer padded to N = 16 (M=128 needs N % 16 == 0)
    // TMEM columns (32-bit): D | ones chunk | A
    static constexpr int D_COL = 0, ONES_COL = H, A_COL = H + 8, USED_COLS = H + 8 + K0 / 2;
    static constexpr int TMEM_COLS = USED_COLS <= 32 ? 32 : USED_COLS <= 64 ? 64 : USED_COLS <= 128 ? 128 : USED_COLS <= 256 ? 256 : 512;
    // weight tiles in the blob (K-major core matrices, chunk-major): bytes
    static __host__ __device__ constexpr int w0_bytes() { return (K0 + 16) * H * 2; }
    static __host__ __device__ constexpr int wh_bytes() { return (H + 16) * H * 2; }
    static __host__ __device__ constexpr int wlast_bytes() { return (H + 16) * NLAST * 2; }

    struct Ctx {
        uint32_t tmem_lane;   // TMEM base address of this thread's lane quadrant (lane field = 32 * (warp % 4))
        uint32_t tmem_base;   // allocation base (lane 0)
        uint32_t sblob;       // shared address of the blob
        uint32_t bar;         // shared address of the MMA-completion mbarrier
        uint32_t phase;
        uint32_t swap;
        tc::ActF af;
    };

    // one thread: issue the K steps of a layer. A chunks at A_COL + 8*kt, then (optionally) the ones chunk.
    static __device__ __forceinline__ void issue_layer(const Ctx& c, uint32_t w_addr, int n, int kt_data, bool bias_chunk) {
        const uint32_t idesc = tc::make_idesc(128, n);
        const uint32_t kstride = uint32_t(n / 8) * 128u, step = 2u * kstride;
        const uint32_t lbo = c.swap ? 128u : kstride, sbo = c.swap ? kstride : 128u;
        for (int kt = 0; kt < kt_data; ++kt)
            tc::mma_ts(c.tmem_base + D_COL, c.tmem_base + A_COL + 8 * kt, tc::make_b_desc(w_addr + kt * step, lbo, sbo), idesc, kt > 0);
        if (bias_chunk)
            tc::mma_ts(c.tmem_base + D_COL, c.tmem_base + ONES_COL, tc::make_b_desc(w_addr + kt_data * step, lbo, sbo), idesc, 1u);
        tc::mma_commit(c.bar);
    }

    // per-CTA set-up after TMEM allocation: the constant ones chunk
    static __device__ __forceinline__ void init_tmem(const Ctx& c) {
        uint32_t ones[8] = {0x00003C00u, 0u, 0u, 0u, 0u, 0u, 0u, 0u};   // (half 1.0, half 0), zeros
        tc::tmem_st8(c.tmem_lane + ONES_COL, ones, 0);
        tc::wait_st();
    }

    // Evaluate the network for the 128 samples of the CTA (row = this thread's layer-0 inputs, build_row(..., ch3_one=true)).
    // Every thread of the CTA must call. Returns false (out undefined) when no thread of the CTA is `active`.
    static __device__ __forceinline__ bool mlp(const NetParams& P, Ctx& c, const uint32_t (&row)[K0 / 2], bool active, float (&out)[4]) {
#pragma unroll
        for (int kt = 0; kt < KT0; ++kt) tc::tmem_st8(c.tmem_lane + A_COL + 8 * kt, row, 8 * kt);
        tc::wait_st();
        tc::fence_before_sync();
        if (!__syncthreads_or(active ? 1 : 0)) return false;
        if (threadIdx.x == 0) {
            tc::fence_after_sync();
            issue_layer(c, c.sblob + P.off_w0, H, KT0, P.has_time != 0);
        }
        const int act = P.activation;
#pragma unroll 1
        for (int l = 0; l <= P.num_hidden; ++l) {
            tc::mbar_wait(c.bar, c.phase); c.phase ^= 1u;
            tc::fence_after_sync();
            // epilogue: accumulator row -> activation -> fp16 A operand of the next layer
#pragma unroll
            for (int c0 = 0; c0 < H; c0 += 32) {
                uint32_t v[32];
                if (H - c0 >= 32) tc::tmem_ld32(c.tmem_lane + D_COL + c0, v); else tc::tmem_ld16(c.tmem_lane + D_COL + c0, v);
                tc::wait_ld();
                uint32_t pk[16];
#pragma unroll
                for (int j = 0; j < 16; ++j) pk[j] = (H - c0 >= 32 || j < 8) ? tc::act_pack(act, c.af, v[2 * j], v[2 * j + 1]) : 0u;
                if (H - c0 >= 32) tc::tmem_st16(c.tmem_lane + A_COL + c0 / 2, pk, 0); else tc::tmem_st8(c.tmem_lane + A_COL + c0 / 2, pk, 0);
            }
            tc::wait_st();
            tc::fence_before_sync();
            __syncthreads();
            if (threadIdx.x == 0) {
                tc::fence_after_sync();
                if (l < P.num_hidden) issue_layer(c, c.sblob + P.off_wh + l * wh_bytes(), H, KTH, true);
                else issue_layer(c, c.sblob + P.off_wlast, NLAST, KTH, true);
            }
        }
        tc::mbar_wait(c.bar, c.phase); c.phase ^= 1u;
        tc::fence_after_sync();
        if (P.output_mode >= D_OUT_RGBO) {
            uint32_t o[4];
            tc::tmem_ld4(c.tmem_lane + D_COL, o);
            tc::wait_ld();
            out[0] = __uint_as_float(o[0]); out[1] = __uint_as_float(o[1]); out[2] = __uint_as_float(o[2]); out[3] = __uint_as_float(o[3]);
        } else {
            uint32_t o;
            tc::tmem_ld1(c.tmem_lane + D_COL, o);
            tc::wait_ld();
            out[0] = __uint_as_float(o); out[1] = out[2] = out[3] = 0.f;
        }
        // the next call's tcgen05.st / mma may only start after every thread has read its outputs
        tc::fence_before_sync();
        return true;
    }
};

struct TcShared {           // fixed part of the dynamic shared memory, placed first
    unsigned long long bar;
    uint32_t tmem_addr;
    int tile;
};

// common prologue: blob -> smem, mbarrier, TMEM allocation
template <class Net>
__device__ __forceinline__ typename Net::Ctx tc_prologue(const NetParams& P, TcShared* sh, uint4* sblob) {
    load_blob(P, sblob);
    const int warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) tc::mbar_init(smem_u32(&sh->bar), 1);
    if (warp == 0) tc::tmem_alloc(smem_u32(&sh->tmem_addr), Net::TMEM_COLS);
    // make the generic-proxy writes of the weights visible to the tensor-core (async) proxy
    asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    typename Net::Ctx c;
    c.tmem_base = sh->tmem_addr;
    c.tmem_lane = c.tmem_base + (uint32_t((warp & 3) * 32) << 16);
    c.sblob = smem_u32(sblob);
    c.bar = smem_u32(&sh->bar);
    c.phase = 0;
    c.swap = uint32_t(P.dbg_flags & 1);
    c.af = tc::make_actf(P.activation, P.act_param);
    Net::init_tmem(c);
    return c;
}
template <class Net>
__device__ __forceinline__ void tc_epilogue(const typename Net::Ctx& c) {
    tc::fence_before_sync();
    __syncthreads();
    if ((threadIdx.x >> 5) == 0) tc::tmem_dealloc(c.tmem_base, Net::TMEM_COLS);
}

template <int H16, int GC16, bool DIR>
__global__ void __launch_bounds__(kTcThreads)
dvr_tc_kernel(const __grid_constant__ NetParams P, const __grid_constant__ DvrParams D) {
    using Net = SrnTc<H16, GC16, DIR>;
    using Core = typename Net::Core;
    extern __shared__ uint4 smem[];
    TcShared* sh = reinterpret_cast<TcShared*>(smem);
    uint4* sblob = smem + 1;
    const int blob16 = (P.blob_bytes + 15) / 16;
    float* stf = reinterpret_cast<float*>(sblob + blob16);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    typename Net::Ctx ctx = tc_prologue<Net>(P, sh, sblob);
    const int tf_floats = D.tf_kind == 0 ? D.tf_R * 5 : (D.tf_kind == 1 ? D.tf_R * 4 : 2);
    if (D.tf_batch == 1)
        for (int i = threadIdx.x; i < tf_floats; i += blockDim.x) stf[i] = __ldg(D.tf + i);
    __syncthreads();

    const float3 bmin = make_float3(P.box_min[0], P.box_min[1], P.box_min[2]);
    const float3 bsize = make_float3(P.box_size[0], P.box_size[1], P.box_size[2]);
    const float3 binv = make_float3(__fdiv_rn(1.f, bsize.x), __fdiv_rn(1.f, bsize.y), __fdiv_rn(1.f, bsize.z));
    // pixel of this thread inside the CTA tile: warps 2x2, lanes 8x4
    const int lx = (warp & 1) * 8 + (lane & 7), ly = (warp >> 1) * 4 + (lane >> 3);

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
        const int x = D.x0 + tx * kTcTileW + lx;
        const int y = D.y0 + ty * kTcTileH + ly;
        const bool in_img = (x < D.x0 + D.tw) && (y < D.y0 + D.th);
        const float* tfp = (D.tf_batch == 1) ? stf : (D.tf + (size_t)b * tf_floats);
        const CamFrame cf = load_camera(D, b);

        float4 pix_color = make_float4(0.f, 0.f, 0.f, 0.f);
        float pix_depth = 0.f;
        int last_steps = 0;
        for (int s = 0; s < D.spp; ++s) {   // CTA-uniform trip count
            const Ray r = make_ray(D, cf, x, y, s, bmin, bsize);
            const float tmax = in_img ? r.tmax : -1.f;
            float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
            float acc_depth = 0.f;
            int steps = 0;
            // lock-step march of the 128 rays, reference renderer_ray_evaluation_stepping_dvr.cuh:84-154
            for (int i = 0;; ++i) {
                const float t = __fmaf_rn(float(i), D.stepsize, r.tmin);
                bool valid = (t <= tmax);
                if (D.early_out) valid = valid && (acc.w < D.alpha_early_out);
                float px = 0.f, py = 0.f, pz = 0.f;
                if (valid) {
                    px = __fmul_rn(__fsub_rn(__fmaf_rn(r.dir.x, t, r.eye.x), bmin.x), binv.x);
                    py = __fmul_rn(__fsub_rn(__fmaf_rn(r.dir.y, t, r.eye.y), bmin.y), binv.y);
                    pz = __fmul_rn(__fsub_rn(__fmaf_rn(r.dir.z, t, r.eye.z), bmin.z), binv.z);
                }
                uint32_t row[Core::K0 / 2];
                Core::build_row(P, ctx.sblob, px, py, pz, r.dir.x, r.dir.y, r.dir.z, true, row);
                float o[4];
                if (!Net::mlp(P, ctx, row, valid, o)) break;
                if (lane == 0) n_exec += 32;
                if (!valid) continue;
                ++steps;
                blend_sample(D, shade_sample(P, D, tfp, o), t, acc, acc_depth);
            }
            n_valid += steps;
            last_steps = steps;
            pix_color.x += acc.x; pix_color.y += acc.y; pix_color.z += acc.z; pix_color.w += acc.w;
            pix_depth += acc_depth * acc.w;
        }
        if (in_img) write_pixel(D, b, x, y, pix_color, pix_depth, last_steps);
    }
    if (D.counters) {
        for (int off = 16; off > 0; off >>= 1) n_valid += __shfl_xor_sync(0xffffffffu, n_valid, off);
        if (lane == 0) { atomicAdd(D.counters + 0, n_valid); atomicAdd(D.counters + 1, n_exec); }
    }
    tc_epilogue<Net>(ctx);
}

template <int H16, int GC16, bool DIR>
__global__ void __launch_bounds__(kTcThreads)
eval_tc_kernel(const __grid_constant__ NetParams P, const __grid_constant__ EvalParams E) {
    using Net = SrnTc<H16, GC16, DIR>;
    using Core = typename Net::Core;
    extern __shared__ uint4 smem[];
    TcShared* sh = reinterpret_cast<TcShared*>(smem);
    uint4* sblob = smem + 1;
    typename Net::Ctx ctx = tc_prologue<Net>(P, sh, sblob);
    const long long n128 = (E.n + 127) / 128;
    for (long long w = blockIdx.x; w < n128; w += gridDim.x) {
        const long long i = w * 128 + threadIdx.x;
        const bool valid = i < E.n;
        float px = 0.f, py = 0.f, pz = 0.f, dx = 0.f, dy = 0.f, dz = 0.f;
        if (valid) {
            px = __ldg(E.pos + 3 * i); py = __ldg(E.pos + 3 * i + 1); pz = __ldg(E.pos + 3 * i + 2);
            if (DIR && E.dir) { dx = __ldg(E.dir + 3 * i); dy = __ldg(E.dir + 3 * i + 1); dz = __ldg(E.dir + 3 * i + 2); }
        }
        uint32_t row[Core::K0 / 2];
        Core::build_row(P, ctx.sblob, px, py, pz, dx, dy, dz, true, row);
        float o[4];
        Net::mlp(P, ctx, row, true, o);
        if (!valid) continue;
        if (P.output_mode == D_OUT_DENSITY) E.out[i] = sigmoidf_(o[0]);
        else if (P.output_mode == D_OUT_DENSITY_DIRECT) E.out[i] = o[0];
        else {
            float4 c;
            if (P.output_mode == D_OUT_RGBO) c = make_float4(sigmoidf_(o[0]), sigmoidf_(o[1]), sigmoidf_(o[2]), softplusf_(o[3]));
            else c = make_float4(fminf(fmaxf(o[0], 0.f), 1.f), fminf(fmaxf(o[1], 0.f), 1.f), fminf(fmaxf(o[2], 0.f), 1.f), fmaxf(o[3], 0.f));
            reinterpret_cast<float4*>(E.out)[i] = c;
        }
    }
    tc_epilogue<Net>(ctx);
}

} // namespace fvsrn
