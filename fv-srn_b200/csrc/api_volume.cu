// C ABI of the ground-truth volume (include/fvsrn_b200.h, "Ground-truth volume"): .cvol container, 3-D texture upload,
// DVR of the density grid through the reference-order marcher (kernels_exact.cuh dvr_grid_kernel), world-space queries.
#include "../../include/fvsrn_b200.h"
#include "launch.cuh"
#include "volume.hpp"
#include <cstring>
#include <mutex>
#include <stdexcept>
#include <string>

using namespace fvsrn;

// shared with api.cu
namespace fvsrn_api {
int fail(const std::string& msg);
DvrParams fill_dvr_params(const fvsrn_dvr_params* p, float* out, bool is_rgbo, const void* any_dev);
}

struct fvsrn_volume {
    Volume vol;
    bool prepared = false;
    int device = -1, sm_count = 0, feature = 0;
    cudaArray_t array = nullptr;
    cudaTextureObject_t tex = 0;
    std::mutex mu;
};

namespace {
#define VOL_TRY try {
#define VOL_CATCH } catch (const std::exception& e) { return fvsrn_api::fail(e.what()); } catch (...) { return fvsrn_api::fail("unknown error"); }
#define VOL_CUDA_OK(expr) do { cudaError_t _e = (expr); if (_e != cudaSuccess) \
    throw std::runtime_error(std::string(#expr) + " failed: " + cudaGetErrorString(_e)); } while (0)

struct Guard {      // run on the volume's device, leave the caller's current device as it was
    int prev = -1; bool switched = false;
    explicit Guard(int dev) { if (dev >= 0 && cudaGetDevice(&prev) == cudaSuccess && prev != dev) switched = cudaSetDevice(dev) == cudaSuccess; }
    ~Guard() { if (switched) cudaSetDevice(prev); }
};

void release(fvsrn_volume* v) {
    Guard g(v->device);
    if (v->tex) cudaDestroyTextureObject(v->tex);
    if (v->array) cudaFreeArray(v->array);
    v->tex = 0; v->array = nullptr; v->prepared = false;
}

GridVolumeParams grid_params(const fvsrn_volume* v, bool unit_box) {
    const VolumeFeature& f = v->vol.features[size_t(v->feature)];
    GridVolumeParams G{};
    G.tex = v->tex; G.rx = int(f.X); G.ry = int(f.Y); G.rz = int(f.Z);
    for (int i = 0; i < 3; ++i) {       // box = +-worldSize/2 (volume_interpolation_grid.cpp:193-198); evaluate() forces the unit cube
        G.box_min[i] = unit_box ? 0.f : float(-double(v->vol.world[i]) / 2.0);
        G.box_size[i] = unit_box ? 1.f : v->vol.world[i];
    }
    return G;
}
} // namespace

extern "C" {

int fvsrn_volume_load_cvol(const void* bytes, size_t n, fvsrn_volume** out) {
    VOL_TRY
    if (!bytes || !out) return fvsrn_api::fail("null argument");
    fvsrn_volume* v = new fvsrn_volume();
    try { v->vol = Volume::load(static_cast<const uint8_t*>(bytes), n); }
    catch (...) { delete v; throw; }
    *out = v;
    return 0;
    VOL_CATCH
}

int fvsrn_volume_save_cvol(const fvsrn_volume* vol, void* buf, size_t cap, size_t* written) {
    VOL_TRY
    if (!vol || !written) return fvsrn_api::fail("null argument");
    const std::vector<uint8_t> b = vol->vol.save();
    *written = b.size();
    if (!buf) return 0;
    if (cap < b.size()) return fvsrn_api::fail("buffer too small");
    std::memcpy(buf, b.data(), b.size());
    return 0;
    VOL_CATCH
}

int fvsrn_volume_from_dense(const float* density_host, int size_x, int size_y, int size_z, const float world_size[3], fvsrn_volume** out) {
    VOL_TRY
    if (!density_host || !world_size || !out) return fvsrn_api::fail("null argument");
    fvsrn_volume* v = new fvsrn_volume();
    try { v->vol = Volume::from_dense(density_host, size_x, size_y, size_z, world_size); }
    catch (...) { delete v; throw; }
    *out = v;
    return 0;
    VOL_CATCH
}

int fvsrn_volume_get_info(const fvsrn_volume* vol, int feature, fvsrn_volume_info* info) {
    if (!vol || !info) return fvsrn_api::fail("null argument");
    if (feature < 0 || feature >= int(vol->vol.features.size())) return fvsrn_api::fail("feature index out of bounds");
    const VolumeFeature& f = vol->vol.features[size_t(feature)];
    std::memset(info, 0, sizeof(*info));
    info->num_features = int(vol->vol.features.size()); info->feature = feature;
    info->size_x = int(f.X); info->size_y = int(f.Y); info->size_z = int(f.Z); info->channels = f.channels; info->data_type = f.type;
    for (int i = 0; i < 3; ++i) {
        info->world_size[i] = vol->vol.world[i];
        info->box_min[i] = float(-double(vol->vol.world[i]) / 2.0); info->box_size[i] = vol->vol.world[i];
    }
    return 0;
}

void fvsrn_volume_free(fvsrn_volume* vol) {
    if (!vol) return;
    release(vol);
    delete vol;
}

int fvsrn_volume_prepare(fvsrn_volume* vol, int device, int feature, void* stream) {
    VOL_TRY
    if (!vol) return fvsrn_api::fail("null volume handle");
    std::lock_guard<std::mutex> lock(vol->mu);
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0) return fvsrn_api::fail("no CUDA device available: the fV-SRN B200 path has no CPU fallback");
    if (device < 0 || device >= count) return fvsrn_api::fail("invalid device index");
    if (feature < 0 || feature >= int(vol->vol.features.size())) return fvsrn_api::fail("feature index out of bounds");
    const VolumeFeature& f = vol->vol.features[size_t(feature)];
    if (f.channels != 1) return fvsrn_api::fail("only 1-channel (density) features are rendered on the B200 path");
    release(vol);
    Guard g(device);
    vol->device = device; vol->feature = feature;
    cudaDeviceProp prop;
    VOL_CUDA_OK(cudaGetDeviceProperties(&prop, device));
    vol->sm_count = prop.multiProcessorCount;
    // MipmapLevel::copyCpuToGpu, volume.cpp:120-166: one array in the feature's own type; integer types are read as normalised
    // floats; unnormalised coordinates, clamp, linear filter
    const int bits = int(f.bytes_per_entry()) * 8;
    const cudaChannelFormatDesc fd = cudaCreateChannelDesc(bits, 0, 0, 0, f.type == VOL_FLOAT ? cudaChannelFormatKindFloat : cudaChannelFormatKindUnsigned);
    VOL_CUDA_OK(cudaMalloc3DArray(&vol->array, &fd, make_cudaExtent(f.X, f.Y, f.Z)));
    cudaMemcpy3DParms cp = {};
    cp.srcPtr = make_cudaPitchedPtr(const_cast<uint8_t*>(f.data.data()), f.X * f.bytes_per_entry(), f.X, f.Y);
    cp.dstArray = vol->array; cp.extent = make_cudaExtent(f.X, f.Y, f.Z); cp.kind = cudaMemcpyHostToDevice;
    VOL_CUDA_OK(cudaMemcpy3D(&cp));
    cudaResourceDesc rd = {}; rd.resType = cudaResourceTypeArray; rd.res.array.array = vol->array;
    cudaTextureDesc td = {};
    td.addressMode[0] = td.addressMode[1] = td.addressMode[2] = cudaAddressModeClamp;
    td.filterMode = cudaFilterModeLinear; td.normalizedCoords = 0;
    td.readMode = f.type == VOL_FLOAT ? cudaReadModeElementType : cudaReadModeNormalizedFloat;
    VOL_CUDA_OK(cudaCreateTextureObject(&vol->tex, &rd, &td, nullptr));
    (void)stream;
    vol->prepared = true;
    return 0;
    VOL_CATCH
}

int fvsrn_render_dvr_grid(fvsrn_volume* vol, const fvsrn_dvr_params* p, float* out_dev, void* stream) {
    VOL_TRY
    if (!vol) return fvsrn_api::fail("null volume handle");
    std::lock_guard<std::mutex> lock(vol->mu);
    if (!vol->prepared) return fvsrn_api::fail("volume is not prepared: call fvsrn_volume_prepare first");
    Guard g(vol->device);
    DvrParams D = fvsrn_api::fill_dvr_params(p, out_dev, false, nullptr);
    if (p->tf_kind == FVSRN_TF_IDENTITY) return fvsrn_api::fail("the grid renderer evaluates piecewise or texture transfer functions");
    if (p->view_times_host) return fvsrn_api::fail("a grid volume has no time keyframes");
    D.b0 = 0; D.nb = p->batch;
    VOL_CUDA_OK(launch_dvr_grid(grid_params(vol, false), D, vol->sm_count, static_cast<cudaStream_t>(stream)));
    return 0;
    VOL_CATCH
}

int fvsrn_evaluate_grid(fvsrn_volume* vol, const float* positions_dev, float* out_dev, int64_t n, void* stream) {
    VOL_TRY
    if (!vol) return fvsrn_api::fail("null volume handle");
    std::lock_guard<std::mutex> lock(vol->mu);
    if (!vol->prepared) return fvsrn_api::fail("volume is not prepared: call fvsrn_volume_prepare first");
    if (n < 0) return fvsrn_api::fail("negative sample count");
    if (n == 0) return 0;
    if (!positions_dev || !out_dev) return fvsrn_api::fail("null device pointer");
    Guard g(vol->device);
    VOL_CUDA_OK(launch_eval_grid(grid_params(vol, true), positions_dev, out_dev, (long long)n, static_cast<cudaStream_t>(stream)));
    return 0;
    VOL_CATCH
}

} // extern "C"
