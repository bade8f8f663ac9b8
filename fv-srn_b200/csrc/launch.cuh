// Launch plumbing shared by api.cu and the per-width instantiation units (inst_hN.cu).
// Replaces the run-time NVRTC assembly of reference renderer/kernel_loader.cpp:197-273 by ahead-of-time
// sm_100a instantiations selected from the network's shape (hidden width, latent channels, direction input).
#pragma once
#include "kernels_tc.cuh"
#ifdef FVSRN_EXPERIMENTS   // measured dead ends kept for reference (SS-mode tcgen05 kernel): not in the product build
#include "kernels_ss.cuh"
#include "kernels_tc8.cuh"
#endif
#include <cstdlib>
#include <cuda_runtime.h>
#include <map>
#include <mutex>
#include <utility>

namespace fvsrn {

struct LaunchCfg {
    int sm_count = 0;
    int blocks_per_sm = 0;   // filled by the launcher (occupancy query)
    int grid = 0;            // filled by the launcher
    int smem = 0;
    int regs = 0;
    int threads = 0;
    int query_only = 0;      // 1: only fill the fields above, do not launch
    long long max_blocks = 0; // upper bound on useful blocks (0 = none; DVR launches derive it from the tile raster)
    int tile_w = 0, tile_h = 0; // DVR kernels: pixels per work-queue tile (filled by the launcher)
};

enum { KIND_DVR = 0, KIND_EVAL = 1, KIND_DVR_TC = 2, KIND_EVAL_TC = 3 };  // _TC: tcgen05/TMEM kernels (kernels_tc.cuh)

// kp points to a DvrParams (KIND_DVR) or EvalParams (KIND_EVAL)
typedef cudaError_t (*launch_fn)(int kind, int gc16, int dir, const NetParams& P, const void* kp, LaunchCfg* cfg, cudaStream_t s);
cudaError_t launch_h2(int, int, int, const NetParams&, const void*, LaunchCfg*, cudaStream_t);
cudaError_t launch_h3(int, int, int, const NetParams&, const void*, LaunchCfg*, cudaStream_t);
cudaError_t launch_h4(int, int, int, const NetParams&, const void*, LaunchCfg*, cudaStream_t);
cudaError_t launch_h6(int, int, int, const NetParams&, const void*, LaunchCfg*, cudaStream_t);
cudaError_t launch_h8(int, int, int, const NetParams&, const void*, LaunchCfg*, cudaStream_t);
// reference-exact DVR kernel (kernels_exact.cuh, inst_exact.cu)
cudaError_t launch_exact(int h16, int gc16, int dir, const NetParams& P, const DvrParams& D, LaunchCfg* cfg, cudaStream_t s);
cudaError_t launch_exact_adjoint(int h16, int gc16, const NetParams& P, const DvrParams& D, LaunchCfg* cfg, cudaStream_t s);   // inst_exact_adj.cu
// ground-truth density grid through the same marcher (kernels_exact.cuh dvr_grid_kernel / eval_grid_kernel)
cudaError_t launch_dvr_grid(const GridVolumeParams& G, const DvrParams& D, int sm_count, cudaStream_t s);
cudaError_t launch_eval_grid(const GridVolumeParams& G, const float* pos, float* out, long long n, cudaStream_t s);
// tensor operators beside the fused path (inst_exact.cu): camera rays, transfer-function evaluation
cudaError_t launch_generate_rays(const float* cam, float fov_y, float aspect, int B, int H, int W, float* start, float* dir, cudaStream_t s);
cudaError_t launch_evaluate_tf(int tf_kind, const float* tf, int R, const float* density, float* out, long long n, float dmin, float dmax,
                               float stepsize, cudaStream_t s);

// CTAs per SM the register file allows: registers live in four 16384-entry partitions (one per SM sub-partition) and a
// warp's registers cannot straddle them, so the limit is per partition: floor(16384 / regs per warp) warps each.
// (A 5-warp CTA at 174 registers: 2 warps per partition -> 8 warps -> ONE CTA, although 2 x 5 x 176 x 32 < 65536.)
inline int tc_blocks_by_regs(int regs_per_thread, int threads) {
    const int per_warp = ((regs_per_thread + 7) / 8 * 8) * 32;
    const int warps_per_partition = 16384 / per_warp;
    return (4 * warps_per_partition) / ((threads + 31) / 32);
}

#ifdef FVSRN_EXPERIMENTS
// SS-mode tcgen05 DVR kernel (kernels_ss.cuh): hidden width 32, density networks without direction input
template <int H16, int GC16, int NS>
cudaError_t launch_ss_n(const NetParams& P, const DvrParams& D0, LaunchCfg* cfg, cudaStream_t s, const void* fn) {
    using Net = SrnSs<H16, GC16, 0, NS>;
    DvrParams D = D0;
    const int tf_floats = D.tf_kind == 0 ? (D.tf_R - 1) * 12 + kTcTfBins / 4 : (D.tf_kind == 1 ? D.tf_R * 4 : 4);
    D.tf_smem_floats = ((tf_floats < 16 ? 16 : tf_floats) + 3) / 4 * 4;
    const int blob = ((P.blob_bytes + 15) / 16) * 16;
    const int smem = int(sizeof(SsShared<NS>)) + blob + D.tf_smem_floats * 4 + Net::ONES_BYTES + NS * Net::A_BYTES;
    cudaError_t e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return e;
    cudaFuncAttributes fa;
    e = cudaFuncGetAttributes(&fa, fn);
    if (e != cudaSuccess) return e;
    const int by_regs = tc_blocks_by_regs(fa.numRegs, kTcThreads);
    const int by_smem = int((227 * 1024) / (size_t(smem) + 1024));
    const int by_tmem = 512 / Net::TMEM_COLS;
    const int bps = max(1, min(min(by_regs, by_smem), by_tmem));
    long long grid = (long long)bps * cfg->sm_count;
    if (cfg->max_blocks > 0 && grid > cfg->max_blocks) grid = cfg->max_blocks;
    if (grid < 1) grid = 1;
    cfg->blocks_per_sm = bps; cfg->grid = int(grid); cfg->smem = smem; cfg->regs = fa.numRegs; cfg->threads = kTcThreads;
    if (cfg->query_only) return cudaSuccess;
    void* args[2] = {const_cast<NetParams*>(&P), &D};
    return cudaLaunchKernel(fn, dim3(unsigned(grid)), dim3(kTcThreads), args, size_t(smem), s);
}
template <int H16, int GC16, bool DIR>
cudaError_t launch_ss(const NetParams& P, const DvrParams& D0, LaunchCfg* cfg, cudaStream_t s, bool* handled) {
    *handled = false;
    if constexpr (H16 == 2 && !DIR) {
        static const int slots = [] { const char* e = std::getenv("FVSRN_SS_SLOTS"); return e ? std::atoi(e) : 0; }();
        if (slots == 0 || P.output_mode >= D_OUT_RGBO) return cudaSuccess;    // default: TS-mode kernel (kernels_tc.cuh)
        const int variant = (P.activation == D_ACT_RELU ? 1 : 0) | (P.fourier_nerf ? 2 : 0);
        *handled = true;
        if constexpr (GC16 == 1) {      // experiment: other slot counts for the paper-default shape only
            if (variant == 3 && slots == 2) return launch_ss_n<H16, GC16, 2>(P, D0, cfg, s, (const void*)dvr_ss_kernel<H16, GC16, 3, 2>);
            if (variant == 3 && slots == 4) return launch_ss_n<H16, GC16, 4>(P, D0, cfg, s, (const void*)dvr_ss_kernel<H16, GC16, 3, 4>);
        }
        const void* fns[4] = {(const void*)dvr_ss_kernel<H16, GC16, 0, 3>, (const void*)dvr_ss_kernel<H16, GC16, 1, 3>,
                              (const void*)dvr_ss_kernel<H16, GC16, 2, 3>, (const void*)dvr_ss_kernel<H16, GC16, 3, 3>};
        return launch_ss_n<H16, GC16, 3>(P, D0, cfg, s, fns[variant]);
    }
    return cudaSuccess;
}
#endif  // FVSRN_EXPERIMENTS

// Per-kernel launch facts (dynamic-smem opt-in done, registers, co-resident CTAs per SM), queried once per
// (kernel, shared-memory size) instead of three runtime-API calls on every launch.
struct KernelFacts { int regs = 0; int bps = 0; };
inline cudaError_t kernel_facts(const void* fn, int threads, int smem, bool is_tc, int tmem_cols, KernelFacts* out) {
    static std::mutex mu;
    static std::map<std::pair<const void*, std::pair<int, int>>, KernelFacts> cache;   // key: kernel, (smem, device)
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    const auto key = std::make_pair(fn, std::make_pair(smem, dev));
    {
        std::lock_guard<std::mutex> lock(mu);
        auto it = cache.find(key);
        if (it != cache.end()) { *out = it->second; return cudaSuccess; }
    }
    e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return e;
    int bps = 0;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, fn, threads, smem);
    if (e != cudaSuccess) return e;
    if (bps < 1) return cudaErrorLaunchOutOfResources;
    cudaFuncAttributes fa;
    e = cudaFuncGetAttributes(&fa, fn);
    if (e != cudaSuccess) return e;
    if (is_tc) {
        // The occupancy API reports 1 CTA/SM for kernels that allocate TMEM; co-residency is in fact bounded by
        // registers, shared memory (+1 KiB reserved per CTA) and the 512 TMEM columns of an SM (measured:
        // profiles/microbench). Surplus CTAs of the persistent grid would only queue, never deadlock.
        const int by_regs = tc_blocks_by_regs(fa.numRegs, threads);
        const int by_smem = int((228 * 1024) / (size_t(smem) + 1024));
        const int by_tmem = 512 / tmem_cols;
        bps = max(1, min(min(by_regs, by_smem), min(by_tmem, 16)));
    }
    KernelFacts f; f.regs = fa.numRegs; f.bps = bps;
    {
        std::lock_guard<std::mutex> lock(mu);
        cache[key] = f;
    }
    *out = f;
    return cudaSuccess;
}

// the tile raster of one DVR launch for a kernel whose CTA (tcgen05 kernels) or warp (mma.sync kernels) owns tile_w x tile_h pixels
inline void set_tiles(DvrParams& d, int tile_w, int tile_h) {
    d.tiles_x = (d.tw + tile_w - 1) / tile_w; d.tiles_y = (d.th + tile_h - 1) / tile_h;
    d.num_tiles = d.tiles_x * d.tiles_y * d.nb;
}

#ifdef FVSRN_EXPERIMENTS
// Measured dead end (A/B on one B200, C3, profiles/r02c_*): the 8-compute-warp, one-sample-per-thread kernel
// (kernels_tc8.cuh; 72 registers, 24 compute warps per SM) renders C3 in 7.19 ms against 5.41 ms for dvr_tc_kernel: the
// march is bound by the latency of its three MMA hand-overs per sample, which only a second sample in flight per
// thread hides - more warps with one sample each do not.
#ifndef FVSRN_TC8
#define FVSRN_TC8 0
#endif
template <int H16, int GC16, bool DIR>
cudaError_t launch_tc8(const NetParams& P, const DvrParams& D0, LaunchCfg* cfg, cudaStream_t s, bool* handled) {
    *handled = false;
    if constexpr (H16 == 2 && !DIR && FVSRN_TC8 != 0) {
        if (P.output_mode >= D_OUT_RGBO) return cudaSuccess;
        *handled = true;
        using Net = SrnTc<H16, GC16, false, 0>;
        DvrParams D = D0;
        set_tiles(D, kTc8TileW, kTc8TileH);
        const int variant = (P.activation == D_ACT_RELU ? 1 : 0) | (P.fourier_nerf ? 2 : 0);
        const void* fns[4] = {(const void*)dvr_tc8_kernel<H16, GC16, 0>, (const void*)dvr_tc8_kernel<H16, GC16, 1>,
                              (const void*)dvr_tc8_kernel<H16, GC16, 2>, (const void*)dvr_tc8_kernel<H16, GC16, 3>};
        const void* fn = fns[variant];
        const int blob = ((P.blob_bytes + 15) / 16) * 16;
        const int smem = int(sizeof(Tc8Shared)) + blob + kTcTfFloats * 4;
        KernelFacts kf;
        const cudaError_t e = kernel_facts(fn, kTc8Threads, smem, true, Net::TMEM_COLS, &kf);
        if (e != cudaSuccess) return e;
        long long grid = (long long)kf.bps * cfg->sm_count;
        if (grid > D.num_tiles) grid = D.num_tiles;
        if (grid < 1) grid = 1;
        cfg->blocks_per_sm = kf.bps; cfg->grid = int(grid); cfg->smem = smem; cfg->regs = kf.regs; cfg->threads = kTc8Threads;
        cfg->tile_w = kTc8TileW; cfg->tile_h = kTc8TileH;
        if (cfg->query_only) return cudaSuccess;
        void* args[2] = {const_cast<NetParams*>(&P), &D};
        return cudaLaunchKernel(fn, dim3(unsigned(grid)), dim3(kTc8Threads), args, size_t(smem), s);
    }
    return cudaSuccess;
}

#endif  // FVSRN_EXPERIMENTS (tc8)

template <int H16, int GC16, bool DIR>
cudaError_t launch_one(int kind, const NetParams& P, const void* kp, LaunchCfg* cfg, cudaStream_t s) {
    using Net = Srn<H16, GC16, DIR>;
#ifdef FVSRN_EXPERIMENTS
    if (kind == KIND_DVR_TC && kp != nullptr) {
        bool handled = false;
        const cudaError_t e = launch_tc8<H16, GC16, DIR>(P, *static_cast<const DvrParams*>(kp), cfg, s, &handled);
        if (e != cudaSuccess || handled) return e;
    }
#endif
#ifdef FVSRN_EXPERIMENTS
    if (kind == KIND_DVR_TC && kp != nullptr) {
        bool handled = false;
        const cudaError_t e = launch_ss<H16, GC16, DIR>(P, *static_cast<const DvrParams*>(kp), cfg, s, &handled);
        if (e != cudaSuccess || handled) return e;
    }
#endif
    const bool is_tc = kind >= KIND_DVR_TC, is_dvr = (kind == KIND_DVR || kind == KIND_DVR_TC);
    const int blob = ((P.blob_bytes + 15) / 16) * 16;
    const int threads = is_tc ? kTcThreads : kWarpsPerBlock * 32;
    const int smem = blob + (is_dvr ? (is_tc ? kTcTfFloats : kTfSmemFloats) * 4 : 0) + (is_tc ? int(sizeof(TcShared)) : kWarpsPerBlock * Net::STAGE_BYTES);
    // tcgen05 kernels: the common cases (ReLU or pre-scaled SnakeAlt, NeRF-style Fourier matrix) get their own instantiations
    constexpr int NV = DIR ? 0 : 2;     // the NeRF recurrence is instantiated for networks without direction input only
    const int variant = (P.activation == D_ACT_RELU ? 1 : (P.activation == tc::D_ACT_SNAKEALT_PRE ? 2 : 0)) + 3 * ((P.fourier_nerf && !DIR) ? 1 : 0);
    const void* tc_dvr[6] = {(const void*)dvr_tc_kernel<H16, GC16, DIR, 0>, (const void*)dvr_tc_kernel<H16, GC16, DIR, 1>, (const void*)dvr_tc_kernel<H16, GC16, DIR, 4>,
                             (const void*)dvr_tc_kernel<H16, GC16, DIR, NV>, (const void*)dvr_tc_kernel<H16, GC16, DIR, NV + 1>, (const void*)dvr_tc_kernel<H16, GC16, DIR, NV + 4>};
    const void* tc_eval[6] = {(const void*)eval_tc_kernel<H16, GC16, DIR, 0>, (const void*)eval_tc_kernel<H16, GC16, DIR, 1>, (const void*)eval_tc_kernel<H16, GC16, DIR, 4>,
                              (const void*)eval_tc_kernel<H16, GC16, DIR, NV>, (const void*)eval_tc_kernel<H16, GC16, DIR, NV + 1>, (const void*)eval_tc_kernel<H16, GC16, DIR, NV + 4>};
    // ... and, for the network shapes of the BASELINE configs (hidden 32 / 64, one 16-channel grid, no direction input), the
    // DVR kernel specialised for the common launch configuration (kernels_tc.cuh TcSpec): SPEC 1 static grid, 2 two keyframes
    const void* tc_dvr_spec = nullptr;
    if constexpr ((H16 == 2 || H16 == 4) && GC16 == 1 && !DIR) {
        if (kind == KIND_DVR_TC && kp != nullptr) {
            const DvrParams& Dq = *static_cast<const DvrParams*>(kp);
            const bool fast = P.output_mode == D_OUT_DENSITY_DIRECT && !P.has_time && Dq.tf_kind == 0 && Dq.tf_batch == 1 && Dq.early_out && Dq.blend_mode == 1 && Dq.spp == 1 && P.grid_encoding == 0;
            if (fast) {
                const void* s1[6] = {(const void*)dvr_tc_kernel<H16, GC16, DIR, 0, 1>, (const void*)dvr_tc_kernel<H16, GC16, DIR, 1, 1>, (const void*)dvr_tc_kernel<H16, GC16, DIR, 4, 1>,
                                     (const void*)dvr_tc_kernel<H16, GC16, DIR, 2, 1>, (const void*)dvr_tc_kernel<H16, GC16, DIR, 3, 1>, (const void*)dvr_tc_kernel<H16, GC16, DIR, 6, 1>};
                const void* s2[6] = {(const void*)dvr_tc_kernel<H16, GC16, DIR, 0, 2>, (const void*)dvr_tc_kernel<H16, GC16, DIR, 1, 2>, (const void*)dvr_tc_kernel<H16, GC16, DIR, 4, 2>,
                                     (const void*)dvr_tc_kernel<H16, GC16, DIR, 2, 2>, (const void*)dvr_tc_kernel<H16, GC16, DIR, 3, 2>, (const void*)dvr_tc_kernel<H16, GC16, DIR, 6, 2>};
                tc_dvr_spec = P.groups[0].frac > 0.f ? s2[variant] : s1[variant];
            }
        }
    }
    const bool normals = kind == KIND_DVR && kp != nullptr && static_cast<const DvrParams*>(kp)->normals != 0;
    // adjoint gradient mode: the pre-activations of the repeated forward pass, per warp (srn_core.cuh Srn::adjoint)
    const int smem_adj = (normals && P.gradient_mode == 2) ? kWarpsPerBlock * Net::pre_words(P.num_hidden) * 4 : 0;
    const void* fn = kind == KIND_DVR ? (normals ? (const void*)dvr_kernel<H16, GC16, DIR, true> : (const void*)dvr_kernel<H16, GC16, DIR, false>)
                   : kind == KIND_EVAL ? (const void*)eval_kernel<H16, GC16, DIR>
                   : kind == KIND_DVR_TC ? (tc_dvr_spec ? tc_dvr_spec : tc_dvr[variant]) : tc_eval[variant];
    KernelFacts kf;
    const cudaError_t e = kernel_facts(fn, threads, smem + smem_adj, is_tc, SrnTc<H16, GC16, DIR, 0>::TMEM_COLS, &kf);
    if (e != cudaSuccess) return e;
    const int bps = kf.bps;
    DvrParams dvr;
    if (is_dvr && kp != nullptr) {
        dvr = *static_cast<const DvrParams*>(kp);
        cfg->tile_w = is_tc ? kTcTileW : kTileW; cfg->tile_h = is_tc ? kTcTileH : kTileH;
        set_tiles(dvr, cfg->tile_w, cfg->tile_h);
        cfg->max_blocks = is_tc ? dvr.num_tiles : (dvr.num_tiles + kWarpsPerBlock - 1) / kWarpsPerBlock;
        kp = &dvr;
    }
    long long grid = (long long)bps * cfg->sm_count;
    if (cfg->max_blocks > 0 && grid > cfg->max_blocks) grid = cfg->max_blocks;
    if (grid < 1) grid = 1;
    cfg->blocks_per_sm = bps; cfg->grid = int(grid); cfg->smem = smem + smem_adj; cfg->regs = kf.regs; cfg->threads = threads;
    if (cfg->query_only) return cudaSuccess;
    if (is_tc) {
        void* args[2] = {const_cast<NetParams*>(&P), const_cast<void*>(kp)};
        return cudaLaunchKernel(fn, dim3(unsigned(grid)), dim3(threads), args, size_t(smem), s);
    }
    switch (kind) {
        case KIND_DVR:
            if (normals) dvr_kernel<H16, GC16, DIR, true><<<int(grid), threads, smem + smem_adj, s>>>(P, *static_cast<const DvrParams*>(kp));
            else dvr_kernel<H16, GC16, DIR, false><<<int(grid), threads, smem, s>>>(P, *static_cast<const DvrParams*>(kp));
            break;
        default: eval_kernel<H16, GC16, DIR><<<int(grid), threads, smem, s>>>(P, *static_cast<const EvalParams*>(kp)); break;
    }
    return cudaGetLastError();
}

template <int H16>
cudaError_t launch_width(int kind, int gc16, int dir, const NetParams& P, const void* kp, LaunchCfg* cfg, cudaStream_t s) {
    if (!dir) {
        switch (gc16) {
            case 0: return launch_one<H16, 0, false>(kind, P, kp, cfg, s);
            case 1: return launch_one<H16, 1, false>(kind, P, kp, cfg, s);
            case 2: return launch_one<H16, 2, false>(kind, P, kp, cfg, s);
        }
    } else {
        switch (gc16) {
            case 0: return launch_one<H16, 0, true>(kind, P, kp, cfg, s);
            case 1: return launch_one<H16, 1, true>(kind, P, kp, cfg, s);
            case 2: return launch_one<H16, 2, true>(kind, P, kp, cfg, s);
        }
    }
    return cudaErrorInvalidValue;
}

} // namespace fvsrn
