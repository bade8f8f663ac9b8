// tcgen05 / TMEM DVR kernel, 8 compute warps per CTA (density networks, hidden width 32): the two M=128 tiles that
// dvr_tc_kernel keeps in flight per CTA as two march steps of the same 128 rays ("slots") are here two HALVES of a
// 16x16 pixel tile - 256 rays, one ray and ONE sample in flight per thread. Same TMEM layout, same weight tiles, same
// MMA warp (it serves the halves alternately), but 8 instead of 4 compute warps per CTA share them, each with half the
// live state: 72 instead of 96 registers, 3 CTAs = 24 compute warps per SM instead of 16. dvr_tc_kernel is bound by
// issue-slot utilisation at 4 compute warps per scheduler (ncu r02a: 63 % issue-active, every warp stalled 85 % of the
// time on its own dependent latencies); 6 warps per scheduler fill more of those slots.
//   warps 0-3: rows 0-7 of the tile (half 0), warps 4-7: rows 8-15 (half 1), warp 8: MMA issue.
// The halves march independently (own mbarriers, own trip counters, own termination flags) and rejoin per tile.
#pragma once
#include "kernels_tc.cuh"

namespace fvsrn {

constexpr int kTc8ComputeThreads = 256;
constexpr int kTc8Threads = 288;
constexpr int kTc8TileW = 16, kTc8TileH = 16;

struct Tc8Shared {          // fixed part of the dynamic shared memory, placed first (64 bytes)
    unsigned long long ready[2];   // half h: operands written, one arrival per compute warp of the half
    unsigned long long done[2];    // half h: MMA chain complete (tcgen05.commit)
    uint32_t tmem_addr;
    int tile;
    int flag[2][2];                // [half][trip & 1] = last trip in which a warp of the half had live rays
    int pad[2];
};

template <int H16, int GC16, int VARIANT>
__global__ void __launch_bounds__(kTc8Threads, 3)
dvr_tc8_kernel(const __grid_constant__ NetParams P, const __grid_constant__ DvrParams D) {
    using Net = SrnTc<H16, GC16, false, VARIANT>;
    using Core = typename Net::Core;
    extern __shared__ uint4 smem[];
    Tc8Shared* sh = reinterpret_cast<Tc8Shared*>(smem);
    uint4* sblob = smem + 4;
    const int blob16 = (P.blob_bytes + 15) / 16;
    float* stf = reinterpret_cast<float*>(sblob + blob16);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const bool is_mma_warp = warp == 8;
    const int half = (warp >> 2) & 1;

    // ---- prologue: weights -> smem, mbarriers, TMEM
    load_blob(P, sblob);
    if (threadIdx.x == 0) {
        sh->flag[0][0] = sh->flag[0][1] = sh->flag[1][0] = sh->flag[1][1] = -1;
        tc::mbar_init(smem_u32(&sh->ready[0]), 4); tc::mbar_init(smem_u32(&sh->ready[1]), 4);
        tc::mbar_init(smem_u32(&sh->done[0]), 1); tc::mbar_init(smem_u32(&sh->done[1]), 1);
    }
    if (is_mma_warp) tc::tmem_alloc(smem_u32(&sh->tmem_addr), Net::TMEM_COLS);
    asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    typename Net::Ctx ctx;
    ctx.tmem_base = sh->tmem_addr;
    ctx.tmem_lane = ctx.tmem_base + (uint32_t((warp & 3) * 32) << 16);
    asm volatile("" : "+r"(ctx.tmem_lane));
    ctx.sblob = smem_u32(sblob);
    ctx.ready = smem_u32(&sh->ready[0]);
    asm volatile("" : "+r"(ctx.ready));
    ctx.done = ctx.ready + 16;
    ctx.phase = 0;
    ctx.af = tc::make_actf(P.activation, P.act_param);
    if (warp < 4) Net::init_tmem(ctx);
    if (D.tf_batch == 1) tc_stage_tf(D, 0, stf);
    __syncthreads();

    const float3 bmin = make_float3(P.box_min[0], P.box_min[1], P.box_min[2]);
    const float3 bsize = make_float3(P.box_size[0], P.box_size[1], P.box_size[2]);
    const float3 binv = make_float3(__fdiv_rn(1.f, bsize.x), __fdiv_rn(1.f, bsize.y), __fdiv_rn(1.f, bsize.z));
    // pixel of this thread inside the CTA tile: half = rows 8h..8h+7; inside a half warps 2x2, lanes 8x4 (as dvr_tc_kernel)
    const int lx = (warp & 1) * 8 + (lane & 7), ly = half * 8 + ((warp >> 1) & 1) * 4 + (lane >> 3);
    const int act = P.activation;

    unsigned long long n_valid = 0, n_exec = 0;
    int trip_id = 0;            // compute warps: trips of their half; MMA warp: trips of half 0
    int trip_id1 = 0;           // MMA warp: trips of half 1
    uint32_t phase1 = 0;        // MMA warp: parity of half 1's ready barrier (ctx.phase is half 0's)
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
            // ------------------------------------------------------------ MMA warp: serves the two halves alternately
            const int last = P.num_hidden - 1;      // density networks: the last layer is the compute warps' dot product
            for (int s = 0; s < D.spp; ++s) {
                const bool any = __syncthreads_or(0);
                bool c0 = any, c1 = any;
                while (c0 | c1) {
#pragma unroll 1
                    for (int l = -1; l <= last; ++l) {
                        if (c0) {
                            tc::mbar_wait(ctx.ready, ctx.phase);
                            tc::fence_after_sync();
                            if (l < 0) Net::issue_layer(ctx, 0, ctx.sblob + P.off_w0, Net::H, Net::KT0, P.has_time != 0);
                            else Net::issue_layer(ctx, 0, ctx.sblob + P.off_wh + l * Net::wh_bytes(), Net::H, Net::KTH, true);
                            ctx.phase ^= 1u;
                        }
                        if (c1) {
                            tc::mbar_wait(ctx.ready + 8, phase1);
                            tc::fence_after_sync();
                            if (l < 0) Net::issue_layer(ctx, 1, ctx.sblob + P.off_w0, Net::H, Net::KT0, P.has_time != 0);
                            else Net::issue_layer(ctx, 1, ctx.sblob + P.off_wh + l * Net::wh_bytes(), Net::H, Net::KTH, true);
                            phase1 ^= 1u;
                        }
                    }
                    // trip i+1 of a half runs iff one of its warps had live rays in trip i (flag written before that warp's
                    // first arrival of trip i, read here after the last ready-wait of trip i)
                    if (c0) { c0 = *reinterpret_cast<volatile int*>(&sh->flag[0][trip_id & 1]) == trip_id; ++trip_id; }
                    if (c1) { c1 = *reinterpret_cast<volatile int*>(&sh->flag[1][trip_id1 & 1]) == trip_id1; ++trip_id1; }
                }
            }
            continue;
        }
        // ---------------------------------------------------------------- compute warps: one ray, one sample per trip
        const int x = D.x0 + tx * kTc8TileW + lx;
        const int y = D.y0 + ty * kTc8TileH + ly;
        const bool in_img = (x < D.x0 + D.tw) && (y < D.y0 + D.th);
        const CamFrame cf = load_camera(D, b);
        const uint32_t done_bar = ctx.done + 8 * half;
        volatile int* my_flags = &sh->flag[half][0];

        float4 pix_color = make_float4(0.f, 0.f, 0.f, 0.f);
        float pix_depth = 0.f;
        int last_steps = 0;
        for (int s = 0; s < D.spp; ++s) {
            const Ray r = make_ray(D, cf, x, y, s, bmin, bsize, D.spp);
            const float tmax = in_img ? r.tmax : -1.f;
            float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
            float acc_depth = 0.f;
            int steps = 0;
            bool cont = true;
            for (int i = 0;; ++i) {
                const float t = rf::fma(D.stepsize, float(i), r.tmin);
                const bool valid = (t <= tmax) && (!D.early_out || (acc.w < D.alpha_early_out));
                if (i == 0) cont = __syncthreads_or(valid ? 1 : 0);
                if (!cont) break;
                const bool warp_active = __any_sync(0xffffffffu, valid);
                if (warp_active && lane == 0) my_flags[trip_id & 1] = trip_id;
                if (warp_active) {
                    float px = 0.f, py = 0.f, pz = 0.f;
                    if (valid) {
                        px = __fmul_rn(__fsub_rn(__fmaf_rn(r.dir.x, t, r.eye.x), bmin.x), binv.x);
                        py = __fmul_rn(__fsub_rn(__fmaf_rn(r.dir.y, t, r.eye.y), bmin.y), binv.y);
                        pz = __fmul_rn(__fsub_rn(__fmaf_rn(r.dir.z, t, r.eye.z), bmin.z), binv.z);
                    }
                    uint32_t row[Core::K0 / 2];
                    Net::fetch_latent(P, px, py, pz, row);
                    Net::build_row(P, px, py, pz, 0.f, 0.f, 0.f, row);
                    Net::store_row(ctx, half, row);
                } else {
                    Net::publish(ctx, half);
                }
                // layer chain of this half: hidden-to-hidden hand-overs, then the last hidden layer + dot product
#pragma unroll 1
                for (int l = 1; l <= P.num_hidden; ++l) {
                    tc::mbar_wait(done_bar, ctx.phase);
                    tc::fence_after_sync();
                    if (warp_active) Net::epilogue(ctx, half, act);
                    Net::publish(ctx, half);
                    ctx.phase ^= 1u;
                }
                tc::mbar_wait(done_bar, ctx.phase);
                tc::fence_after_sync();
                float o[4];
                o[0] = warp_active ? Net::epilogue_dot(P, ctx, half, act) : 0.f;
                o[1] = o[2] = o[3] = 0.f;
                ctx.phase ^= 1u;
                tc::fence_before_sync();
                cont = my_flags[trip_id & 1] == trip_id;
                ++trip_id;
                if (lane == 0) n_exec += 32;
                const float4 c = tc_shade_sample(P, D, stf, o);
                steps += valid ? 1 : 0;
                blend_sample_pred(D, D.blend_mode, valid, c, t, acc, acc_depth);
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
