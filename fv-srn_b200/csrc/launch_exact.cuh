// Launch helper of the reference-exact DVR kernel, shared by its two translation units (inst_exact.cu, inst_exact_adj.cu).
#pragma once
#include "launch.cuh"
#include "kernels_exact.cuh"
#include <algorithm>
namespace fvsrn {

template <int H16, int GC16 = 1, bool DIR = false, bool ADJ = false>
static cudaError_t launch_exact_h(const NetParams& P, const DvrParams& D0, LaunchCfg* cfg, cudaStream_t s) {
    using Core = Srn<H16, GC16, DIR>;
    DvrParams D = D0;
    set_tiles(D, kTileW, kTileH);
    cfg->tile_w = kTileW; cfg->tile_h = kTileH;
    cfg->max_blocks = (D.num_tiles + kWarpsPerBlock - 1) / kWarpsPerBlock;
    const int blob = ((P.blob_bytes + 15) / 16) * 16;
    const int threads = kWarpsPerBlock * 32;
    const int smem = blob + kTfSmemFloats * 4 + kWarpsPerBlock * Core::STAGE_BYTES;
    const void* fn = (const void*)dvr_exact_kernel<H16, GC16, DIR, ADJ>;
    KernelFacts kf;
    const cudaError_t e = kernel_facts(fn, threads, smem, false, 512, &kf);
    if (e != cudaSuccess) return e;
    long long grid = (long long)kf.bps * cfg->sm_count;
    if (cfg->max_blocks > 0 && grid > cfg->max_blocks) grid = cfg->max_blocks;
    if (grid < 1) grid = 1;
    cfg->blocks_per_sm = kf.bps; cfg->grid = int(grid); cfg->smem = smem; cfg->regs = kf.regs; cfg->threads = threads;
    if (cfg->query_only) return cudaSuccess;
    dvr_exact_kernel<H16, GC16, DIR, ADJ><<<int(grid), threads, smem, s>>>(P, D);
    return cudaGetLastError();
}

} // namespace fvsrn
