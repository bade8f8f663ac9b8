// AOT sm_100a instantiations of the DVR and batch-evaluation kernels for hidden width 48.
#include "launch.cuh"
namespace fvsrn {
cudaError_t launch_h3(int kind, int gc16, int dir, const NetParams& P, const void* kp, LaunchCfg* cfg, cudaStream_t s) {
    return launch_width<3>(kind, gc16, dir, P, kp, cfg, s);
}
} // namespace fvsrn
