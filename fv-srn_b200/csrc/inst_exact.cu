// AOT sm_100a instantiations of the reference-exact DVR kernel (kernels_exact.cuh). This unit is compiled with
// --use_fast_math, like the reference compiles its kernels (kernel_loader.cpp:229-241), so that the header-implemented
// half functions it shares with the reference (hsin, hcos, __hdiv of cuda_fp16.hpp) expand to the same instructions.
#include "launch.cuh"
#include "kernels_exact.cuh"
namespace fvsrn {

template <int H16>
static cudaError_t launch_exact_h(const NetParams& P, const DvrParams& D0, LaunchCfg* cfg, cudaStream_t s) {
    using Core = Srn<H16, 1, false>;
    DvrParams D = D0;
    set_tiles(D, kTileW, kTileH);
    cfg->tile_w = kTileW; cfg->tile_h = kTileH;
    cfg->max_blocks = (D.num_tiles + kWarpsPerBlock - 1) / kWarpsPerBlock;
    const int blob = ((P.blob_bytes + 15) / 16) * 16;
    const int threads = kWarpsPerBlock * 32;
    const int smem = blob + kTfSmemFloats * 4 + kWarpsPerBlock * Core::STAGE_BYTES;
    const void* fn = (const void*)dvr_exact_kernel<H16>;
    KernelFacts kf;
    const cudaError_t e = kernel_facts(fn, threads, smem, false, 512, &kf);
    if (e != cudaSuccess) return e;
    long long grid = (long long)kf.bps * cfg->sm_count;
    if (cfg->max_blocks > 0 && grid > cfg->max_blocks) grid = cfg->max_blocks;
    if (grid < 1) grid = 1;
    cfg->blocks_per_sm = kf.bps; cfg->grid = int(grid); cfg->smem = smem; cfg->regs = kf.regs; cfg->threads = threads;
    if (cfg->query_only) return cudaSuccess;
    dvr_exact_kernel<H16><<<int(grid), threads, smem, s>>>(P, D);
    return cudaGetLastError();
}

cudaError_t launch_exact(int h16, const NetParams& P, const DvrParams& D, LaunchCfg* cfg, cudaStream_t s) {
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
