// AOT sm_100a instantiations of the reference-exact DVR kernel WITH the adjoint pass (GRADIENT_MODE 2, kernels_exact.cuh
// SrnExact::adjoint), compiled with --use_fast_math like inst_exact.cu. Frames in adjoint gradient mode run here.
#define FVSRN_EXACT_WITHOUT_GRID_KERNELS
#include "launch_exact.cuh"
namespace fvsrn {

cudaError_t launch_exact_adjoint(int h16, int gc16, const NetParams& P, const DvrParams& D, LaunchCfg* cfg, cudaStream_t s) {
    if (gc16 == 2) {
        if (h16 == 2) return launch_exact_h<2, 2, false, true>(P, D, cfg, s);
        if (h16 == 4) return launch_exact_h<4, 2, false, true>(P, D, cfg, s);
        return cudaErrorInvalidValue;
    }
    if (gc16 != 1) return cudaErrorInvalidValue;
    switch (h16) {
        case 2: return launch_exact_h<2, 1, false, true>(P, D, cfg, s);
        case 3: return launch_exact_h<3, 1, false, true>(P, D, cfg, s);
        case 4: return launch_exact_h<4, 1, false, true>(P, D, cfg, s);
        case 6: return launch_exact_h<6, 1, false, true>(P, D, cfg, s);
        case 8: return launch_exact_h<8, 1, false, true>(P, D, cfg, s);
    }
    return cudaErrorInvalidValue;
}

} // namespace fvsrn
