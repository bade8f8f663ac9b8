// C ABI of libfvsrn_b200.so (see include/fvsrn_b200.h). Host logic mirrors the reference's module code:
// SceneNetwork builders / .volnet I/O (renderer/volume_interpolation_network.cpp), constant packing + launch
// (image_evaluator_simple.cpp:198-361, volume_interpolation.cpp:26-127). No CPU fallback: compute entry points
// need a CUDA device.
#include "../../include/fvsrn_b200.h"
#include "launch.cuh"
#include "network.hpp"
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <mutex>
#include <string>
#include <vector>

using namespace fvsrn;

namespace {
thread_local std::string g_last_error;

int fail(const std::string& msg) { g_last_error = msg; return 1; }

#define FV_TRY try {
#define FV_CATCH } catch (const std::exception& e) { return fail(e.what()); } catch (...) { return fail("unknown error"); }
#define CUDA_OK(expr) do { cudaError_t _e = (expr); if (_e != cudaSuccess) \
    throw std::runtime_error(std::string(#expr) + " failed: " + cudaGetErrorString(_e)); } while (0)

// Entry points run on the network's device and leave the caller's current device as they found it.
struct DeviceGuard {
    int prev = -1; bool switched = false;
    explicit DeviceGuard(int dev) {
        if (dev < 0) return;
        if (cudaGetDevice(&prev) != cudaSuccess) { prev = -1; return; }
        if (prev != dev) { switched = cudaSetDevice(dev) == cudaSuccess; }
    }
    ~DeviceGuard() { if (switched && prev >= 0) cudaSetDevice(prev); }
    DeviceGuard(const DeviceGuard&) = delete;
    DeviceGuard& operator=(const DeviceGuard&) = delete;
};

struct DeviceGrid {
    void* data = nullptr; int C = 0, Z = 0, Y = 0, X = 0;     // channel-last fp16 (or bytes, for byte-encoded grids): mma.sync path
    int encoding = 0; std::vector<float> offset, scale;       // LatentGrid::Encoding and its per-channel constants
    std::vector<cudaArray_t> arrays;                          // C/4 fp16 RGBA 3-D arrays + texture objects (tcgen05 path)
    std::vector<cudaTextureObject_t> tex;
    std::vector<cudaArray_t> arrays32;                        // the same (fp16-rounded) texels as fp32 RGBA textures, created on
    std::vector<cudaTextureObject_t> tex32;                   // first use of the reference-exact mode (the reference's own format)
    // blended-keyframe scratch grids only (fvsrn_network::d_blend): surfaces over `arrays`, what the grid currently holds,
    // and the events after the last kernels that read it ([0] on the caller's stream, [1] on the handle's internal stream)
    std::vector<cudaSurfaceObject_t> surf;
    int key_lo = -1, key_hi = -1; float key_frac = -1.f;
    cudaEvent_t ev_read[2] = {nullptr, nullptr};
};
} // namespace

struct fvsrn_network {
    Network net;
    // GPU resources (valid when prepared)
    bool prepared = false;
    int device = -1, sm_count = 0;
    uint4* d_blob = nullptr; int blob_bytes = 0;
    int blob_bytes_adjoint = 0, off_adj_wh = 0, off_adj_w0f = 0, off_adj_w0g = 0;   // blob + transposed weights of the adjoint pass
    int off_w0 = 0, off_wh = 0, off_wlast = 0, off_bias = 0, off_fourier = 0;
    // tcgen05 path: K-major UMMA weight tiles (+ bias chunks) and the Fourier matrix
    uint4* d_tc_blob = nullptr; int tc_blob_bytes = 0;
    int tc_off_w0 = 0, tc_off_wh = 0, tc_off_wlast = 0, tc_off_fourier = 0;
    int tc_activation = 0;      // activation code the tcgen05 kernels run (SnakeAlt may become its pre-scaled form)
    int kernel_mode = FVSRN_KERNEL_TCGEN05;
    int gradient_mode = FVSRN_GRADIENT_OFF_OR_DIRECT;      // VolumeInterpolationNetwork::gradientMode_ (a module setting, not part of .volnet)
    float fd_stepsize = 1.f / 256.f;                       // finiteDifferencesStepsize_, volume_interpolation_network.cpp:1423
    float adjoint_grid_step = 0.f;                         // adjoint mode: central-difference step through the latent grid; 0 = 1 / (4 * resolution), :1424, 1809-1812
    int off_wlast_rows = 0;
    int H16 = 0, GC16 = 0, dir = 0, num_hidden = 0, activation = 0; float act_param = 1.f;
    std::vector<DeviceGrid> d_time, d_ens;
    // tcgen05 path, time-interpolated grids: the two keyframes of a launch are blended ONCE into one of these scratch grids
    // (texels = fp16(a + frac (b - a))) and the kernels then run their static-grid form - the temporal lerp is linear and
    // commutes with the trilinear filter, so every sample saves the second set of fetches (see blend_keyframes)
    std::vector<DeviceGrid> d_blend; int blend_next = 0;
    unsigned int* d_tile_counter = nullptr;   // ring of work-queue counters (one per launch in flight)
    int tile_counter_slots = 0, tile_counter_next = 0;
    float* d_scratch = nullptr; size_t scratch_bytes = 0;  // staging for the *_host entry points
    unsigned long long launches = 0;
    cudaStream_t aux_stream = nullptr;          // multi-view launches alternate between the caller's stream and this one, so
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;   // that the next view's CTAs fill the SMs the previous view drains
    // A handle is used by one host thread at a time (as the reference's modules); concurrent calls on one
    // handle are serialised here because the work-queue counter ring and the staging buffer are per-handle state.
    std::mutex mu;
};

namespace {

void release_gpu(fvsrn_network* h) {
    DeviceGuard guard(h->device);
    if (h->d_blob) cudaFree(h->d_blob);
    if (h->d_tc_blob) cudaFree(h->d_tc_blob);
    h->d_tc_blob = nullptr;
    for (auto* v : {&h->d_time, &h->d_ens, &h->d_blend})
        for (auto& g : *v) {
            if (g.data) cudaFree(g.data);
            for (auto sf : g.surf) cudaDestroySurfaceObject(sf);
            for (auto e : g.ev_read) if (e) cudaEventDestroy(e);
            for (auto t : g.tex) cudaDestroyTextureObject(t);
            for (auto a : g.arrays) cudaFreeArray(a);
            if (!g.encoding) for (auto t : g.tex32) cudaDestroyTextureObject(t);
            for (auto a : g.arrays32) cudaFreeArray(a);
        }
    if (h->d_tile_counter) cudaFree(h->d_tile_counter);
    if (h->d_scratch) cudaFree(h->d_scratch);
    if (h->aux_stream) cudaStreamDestroy(h->aux_stream);
    if (h->ev_fork) cudaEventDestroy(h->ev_fork);
    if (h->ev_join) cudaEventDestroy(h->ev_join);
    h->aux_stream = nullptr; h->ev_fork = h->ev_join = nullptr;
    h->d_blob = nullptr; h->d_time.clear(); h->d_ens.clear(); h->d_blend.clear(); h->blend_next = 0; h->d_tile_counter = nullptr;
    h->d_scratch = nullptr; h->scratch_bytes = 0; h->prepared = false;
}

// fp16 channel-last [Z][Y][X][C] copy of a latent grid (decodes byte encodings on the way)
std::vector<uint16_t> grid_channel_last(const LatentGrid& g) {
    std::vector<uint16_t> out(size_t(g.C) * g.Z * g.Y * g.X);
    for (int z = 0; z < g.Z; ++z) for (int y = 0; y < g.Y; ++y) for (int x = 0; x < g.X; ++x) {
        uint16_t* o = out.data() + ((size_t(z) * g.Y + y) * g.X + x) * g.C;
        for (int c = 0; c < g.C; ++c) o[c] = f32_to_f16(g.decoded(c, z, y, x));
    }
    return out;
}

// Fragment-ordered constant blob (see srn_core.cuh): W0 | hidden W | last W | biases | Fourier matrix
void build_blob(fvsrn_network* h, std::vector<uint8_t>& blob) {
    const Network& n = h->net;
    std::string why;
    if (!n.valid(&why)) throw std::runtime_error("scene network is not valid: " + why);
    if (n.num_fourier <= 0) throw std::runtime_error("unsupported on the B200 path: networks without Fourier features");
    const int H = n.input_channels_out();
    const int GC = n.latent_channels();
    if (!(H == 32 || H == 48 || H == 64 || H == 96 || H == 128))
        throw std::runtime_error("unsupported hidden width " + std::to_string(H) + " (supported: 32, 48, 64, 96, 128)");
    if (!(GC == 0 || GC == 16 || GC == 32)) throw std::runtime_error("unsupported latent channel count " + std::to_string(GC));
    const int L = int(n.layers.size());
    if (L < 2) throw std::runtime_error("need at least one hidden layer and the output layer");
    const Layer& l0 = n.layers[0];
    if (l0.cout != H) throw std::runtime_error("hidden width must equal 4(+4)+2*numFourier (reference :142-152)");
    for (int i = 1; i + 1 < L; ++i)
        if (n.layers[i].cin != H || n.layers[i].cout != H) throw std::runtime_error("all hidden layers must have the same width");
    const Layer& last = n.layers[L - 1];
    if (last.cin != H || last.cout > 8) throw std::runtime_error("bad output layer");
    const int act = l0.activation;
    for (int i = 0; i + 1 < L; ++i) {
        if (n.layers[i].activation != act) throw std::runtime_error("all hidden layers must share one activation");
        if (n.layers[i].activation_param != l0.activation_param) throw std::runtime_error("Extra parameter of the activation must be the same over all layers");
    }
    if (act == ACT_SIGMOID) throw std::runtime_error("Sigmoid is not a valid hidden activation");
    h->H16 = H / 16; h->GC16 = GC / 16; h->dir = n.has_direction ? 1 : 0;
    h->num_hidden = L - 2; h->activation = act; h->act_param = l0.activation_param;

    auto frag_matrix = [&](auto&& at, int K, int N, int Npad, std::vector<uint8_t>& dst) {
        // B fragment of mma.m16n8k16 (col-major B[k][n] = at(n, k)): lane (g = lane/4, t = lane%4) holds
        // b0 = at(n0+g, k0+2t..2t+1), b1 = at(n0+g, k0+2t+8..2t+9)
        const int KT = K / 16, NT = Npad / 8;
        auto W = [&](int nn, int kk) -> uint16_t { return (nn < N) ? at(nn, kk) : uint16_t(0); };
        for (int kt = 0; kt < KT; ++kt)
            for (int ntp = 0; ntp < (NT + 1) / 2; ++ntp)
                for (int lane = 0; lane < 32; ++lane)
                    for (int j = 0; j < (NT >= 2 ? 2 : 1); ++j) {
                        const int nt = 2 * ntp + j, g = lane / 4, t = lane % 4;
                        const int nn = nt * 8 + g, k0 = kt * 16 + 2 * t;
                        const uint16_t v[4] = {W(nn, k0), W(nn, k0 + 1), W(nn, k0 + 8), W(nn, k0 + 9)};
                        const uint8_t* p = reinterpret_cast<const uint8_t*>(v);
                        dst.insert(dst.end(), p, p + 8);
                    }
    };
    auto frag_layer = [&](const Layer& l, int K, int N, int Npad, std::vector<uint8_t>& dst) {
        frag_matrix([&](int nn, int kk) { return l.w(nn, kk); }, K, N, Npad, dst);
    };
    blob.clear();
    h->off_w0 = 0;
    frag_layer(l0, H + GC, H, H, blob);
    h->off_wh = int(blob.size());
    for (int i = 1; i + 1 < L; ++i) frag_layer(n.layers[i], H, H, H, blob);
    h->off_wlast = int(blob.size());
    frag_layer(last, H, last.cout, 8, blob);
    h->off_bias = int(blob.size());
    for (int i = 0; i + 1 < L; ++i) {
        const uint8_t* p = reinterpret_cast<const uint8_t*>(n.layers[i].bias.data());
        blob.insert(blob.end(), p, p + 2 * H);
    }
    for (int i = 0; i < 8; ++i) {
        const uint16_t v = i < last.cout ? last.bias[i] : uint16_t(0);
        blob.push_back(uint8_t(v & 0xff)); blob.push_back(uint8_t(v >> 8));
    }
    while (blob.size() % 16) blob.push_back(0);
    h->off_fourier = int(blob.size());
    {
        const uint8_t* p = reinterpret_cast<const uint8_t*>(n.fourier.data());
        blob.insert(blob.end(), p, p + 2 * n.fourier.size());
    }
    while (blob.size() % 16) blob.push_back(0);
    // the last layer once more as plain rows [cin][8 outputs] for the sequential fp16 chains of the reference-exact mode
    // (4 / 6-output last layers, renderer_volume_tensorcores.cuh:1045-1052, 1106-1115)
    h->off_wlast_rows = int(blob.size());
    for (int k = 0; k < H; ++k)
        for (int o = 0; o < 8; ++o) {
            const uint16_t v = o < last.cout ? last.w(o, k) : uint16_t(0);
            blob.push_back(uint8_t(v & 0xff)); blob.push_back(uint8_t(v >> 8));
        }
    h->blob_bytes = int(blob.size());
    // Adjoint pass (GRADIENT_MODE 2, renderer_volume_tensorcores.cuh:1205-1530): the same layers with the roles of inputs and
    // outputs exchanged, adj_in[cin] = sum_cout W[cout][cin] adj_out[cout] - hidden layers, then layer 0 split into its first H
    // inputs (position, time, Fourier features) and its 16-channel latent groups. Kept behind blob_bytes: only the kernel that
    // runs the adjoint pass copies them to shared memory.
    h->off_adj_wh = int(blob.size());
    for (int i = 1; i + 1 < L; ++i) {
        const Layer& l = n.layers[i];
        frag_matrix([&](int cin, int cout) { return l.w(cout, cin); }, H, H, H, blob);
    }
    h->off_adj_w0f = int(blob.size());
    frag_matrix([&](int cin, int cout) { return l0.w(cout, cin); }, H, H, H, blob);
    h->off_adj_w0g = int(blob.size());
    for (int g = 0; g < GC / 16; ++g)
        frag_matrix([&](int c, int cout) { return l0.w(cout, H + 16 * g + c); }, H, 16, 16, blob);
    h->blob_bytes_adjoint = int(blob.size());
}

// Blob of the tcgen05 path (see srn_tc.cuh / kernels_tc.cuh): per layer the B operand W[n][k] as K-major, no-swizzle
// UMMA tiles. A K=16 instruction step reads two K chunks of 8; a chunk holds N/8 core matrices (8 rows n x 16 B of k)
// back to back, so LBO (K direction) = (N/8)*128 B and SBO (N direction) = 128 B. After the data steps every layer
// carries one more step whose k=0 column is the bias (multiplied by the constant "ones" A chunk). Layer 0 of a network
// without time input has its bias in input column 3 instead (the row builder writes 1.0 there).
void build_tc_blob(fvsrn_network* h, std::vector<uint8_t>& blob) {
    const Network& n = h->net;
    const int H = 16 * h->H16, GC = 16 * h->GC16, L = int(n.layers.size());
    // SnakeAlt evaluates cos(q*v), q = 2*param. When q is a power of two the weights and biases of every activated layer
    // are multiplied by q here - exact in fp16 - so that the accumulators hold q*v and the epilogue saves the multiply
    // (srn_tc.cuh, D_ACT_SNAKEALT_PRE).
    float pre = 1.f;
    h->tc_activation = h->activation;
    if (h->activation == ACT_SNAKEALT) {
        const float q = 2.f * h->act_param;
        int e = 0;
        bool ok = q >= 1.f && q <= 16.f && std::frexp(q, &e) == 0.5f;
        for (int i = 0; ok && i + 1 < L; ++i) {
            for (uint16_t w : n.layers[i].weights) ok = ok && std::fabs(f16_to_f32(w) * q) < 60000.f;
            for (uint16_t b : n.layers[i].bias) ok = ok && std::fabs(f16_to_f32(b) * q) < 60000.f;
        }
        if (ok) { pre = q; h->tc_activation = 6; }
    }
    auto put_layer = [&](const Layer& l, int K, int N, int Npad, bool bias_in_col3, float scale) {
        const size_t base = blob.size();
        blob.resize(base + size_t(K + 16) * Npad * 2, 0);
        uint16_t* dst = reinterpret_cast<uint16_t*>(blob.data() + base);
        auto at = [&](int nn, int kk) -> uint16_t& { return dst[size_t(kk / 8) * (Npad / 8) * 64 + size_t(nn / 8) * 64 + (nn % 8) * 8 + (kk % 8)]; };
        auto sc = [&](uint16_t v) -> uint16_t { return scale == 1.f ? v : f32_to_f16(f16_to_f32(v) * scale); };
        for (int nn = 0; nn < N; ++nn) {
            for (int kk = 0; kk < K; ++kk) at(nn, kk) = sc(l.w(nn, kk));
            if (bias_in_col3) at(nn, 3) = sc(l.bias[nn]);
            else at(nn, K) = sc(l.bias[nn]);
        }
    };
    blob.clear();
    h->tc_off_w0 = 0;
    put_layer(n.layers[0], H + GC, H, H, !n.has_time, pre);
    h->tc_off_wh = int(blob.size());
    for (int i = 1; i + 1 < L; ++i) put_layer(n.layers[i], H, H, H, false, pre);
    h->tc_off_wlast = int(blob.size());
    put_layer(n.layers[L - 1], H, n.layers[L - 1].cout, 16, false, 1.f);
    h->tc_off_fourier = int(blob.size());
    {
        const uint8_t* p = reinterpret_cast<const uint8_t*>(n.fourier.data());
        blob.insert(blob.end(), p, p + 2 * n.fourier.size());
    }
    while (blob.size() % 16) blob.push_back(0);
    h->tc_blob_bytes = int(blob.size());
}

launch_fn pick_launcher(int H16) {
#ifdef FVSRN_ONLY_H2     // kernel-tuning builds (make FAST=1)
    return H16 == 2 ? launch_h2 : nullptr;
#else
    switch (H16) {
        case 2: return launch_h2; case 3: return launch_h3; case 4: return launch_h4;
        case 6: return launch_h6; case 8: return launch_h8;
    }
    return nullptr;
#endif
}

// fp32 RGBA textures of one latent grid for the reference-exact mode: LatentGrid::GPUArray as the reference creates it
// (volume_interpolation_network.cpp:470-504), holding the fp16-rounded voxel values every kernel family of this
// library sees
void ensure_tex32(DeviceGrid& d, const LatentGrid& g) {
    if (!d.tex32.empty()) return;
    if (g.encoding != GRID_FLOAT) { d.tex32 = d.tex; return; }      // byte grids: the unorm8 textures are the reference's format already
    std::vector<float> sub(size_t(g.Z) * g.Y * g.X * 4);
    for (int q = 0; q < g.C / 4; ++q) {
        for (int z = 0; z < g.Z; ++z) for (int y = 0; y < g.Y; ++y) for (int x = 0; x < g.X; ++x)
            for (int c = 0; c < 4; ++c)
                sub[((size_t(z) * g.Y + y) * g.X + x) * 4 + c] = f16_to_f32(f32_to_f16(g.decoded(q * 4 + c, z, y, x)));
        cudaChannelFormatDesc fd = cudaCreateChannelDesc<float4>();
        cudaArray_t arr = nullptr;
        CUDA_OK(cudaMalloc3DArray(&arr, &fd, make_cudaExtent(g.X, g.Y, g.Z)));
        d.arrays32.push_back(arr);
        cudaMemcpy3DParms cp = {};
        cp.srcPtr = make_cudaPitchedPtr(sub.data(), size_t(g.X) * 16, g.X, g.Y);
        cp.dstArray = arr; cp.extent = make_cudaExtent(g.X, g.Y, g.Z); cp.kind = cudaMemcpyHostToDevice;
        CUDA_OK(cudaMemcpy3D(&cp));
        cudaResourceDesc rd = {}; rd.resType = cudaResourceTypeArray; rd.res.array.array = arr;
        cudaTextureDesc td = {};
        td.addressMode[0] = td.addressMode[1] = td.addressMode[2] = cudaAddressModeClamp;
        td.filterMode = cudaFilterModeLinear; td.readMode = cudaReadModeElementType; td.normalizedCoords = 1;
        cudaTextureObject_t t = 0;
        CUDA_OK(cudaCreateTextureObject(&t, &rd, &td, nullptr));
        d.tex32.push_back(t);
    }
}

// channel group i (16 channels) of a channel-last device grid: fp16 halves or bytes
const __half* grid_ptr(const DeviceGrid& d, int i) {
    return d.encoding ? reinterpret_cast<const __half*>(static_cast<const uint8_t*>(d.data) + 16 * i)
                      : static_cast<const __half*>(d.data) + 16 * i;
}
// EncodeGridValue constants of latent-channel group g. The reference decodes keyframe B with A's offset / scale
// (renderer_volume_tensorcores.cuh:586-587, EncodeGridValue<.., false> on both) - a slip that only matters when the
// keyframes were encoded with different ranges; B's own constants are used here (the reference-exact kernel, which is
// compared bit for bit with the reference kernel, follows the reference and ignores enc_*_b).
void fill_encoding(NetParams& P, int g, const DeviceGrid& A, const DeviceGrid& B, int i) {
    P.grid_encoding = A.encoding;
    if (!A.encoding) return;
    if (A.encoding != B.encoding) throw std::runtime_error("all latent grids must share one encoding");
    for (int c = 0; c < 16; ++c) {
        P.enc_offset_a[16 * g + c] = A.offset[16 * i + c]; P.enc_scale_a[16 * g + c] = A.scale[16 * i + c];
        P.enc_offset_b[16 * g + c] = B.offset[16 * i + c]; P.enc_scale_b[16 * g + c] = B.scale[16 * i + c];
    }
}

// device-side view of the network for a given absolute time / ensemble (reference fillConstantMemory :1290-1352)
NetParams make_net_params(const fvsrn_network* h, float time_abs, int ensemble, bool unit_box, bool tc) {
    const Network& n = h->net;
    NetParams P{};
    P.gradient_mode = h->gradient_mode; P.fd_stepsize = h->fd_stepsize; P.off_wlast_rows = h->off_wlast_rows;
    if (h->gradient_mode == FVSRN_GRADIENT_ADJOINT_METHOD) {
        // VolumeInterpolationNetwork::fillConstantMemory, :1809-1812: stepsize = 1 / (float(gridResolution) * scale), scale 4 by default;
        // it fills both finiteDifferencesStepsize and latentGridDifferencesStepSize (:1409-1410)
        int res = 1;
        if (n.has_latent_grid) res = !n.time_grids.empty() ? n.time_grids[0].X : n.ensemble_grids[0].X;
        P.fd_stepsize = h->adjoint_grid_step > 0.f ? h->adjoint_grid_step : 1.0f / (float(res) * 4.0f);
    }
    if (tc) {
        P.blob = h->d_tc_blob; P.blob_bytes = h->tc_blob_bytes;
        P.off_w0 = h->tc_off_w0; P.off_wh = h->tc_off_wh; P.off_wlast = h->tc_off_wlast; P.off_bias = 0; P.off_fourier = h->tc_off_fourier;
    } else {
        P.blob = h->d_blob; P.blob_bytes = h->gradient_mode == FVSRN_GRADIENT_ADJOINT_METHOD ? h->blob_bytes_adjoint : h->blob_bytes;
        P.off_adj_wh = h->off_adj_wh; P.off_adj_w0f = h->off_adj_w0f; P.off_adj_w0g = h->off_adj_w0g;
        P.off_w0 = h->off_w0; P.off_wh = h->off_wh; P.off_wlast = h->off_wlast; P.off_bias = h->off_bias; P.off_fourier = h->off_fourier;
    }
    P.num_hidden = h->num_hidden; P.act_param = h->act_param;
    P.activation = tc ? h->tc_activation : h->activation;
    P.output_mode = n.output_mode; P.has_time = n.has_time ? 1 : 0; P.dir_in_fourier = n.direction_in_fourier ? 1 : 0;
    P.time_value = 0.f;
    int g = 0;
    if (n.has_latent_grid) {
        if (!h->d_time.empty()) {
            const float t = n.interpolate_time(time_abs);
            const int lo = std::min(int(t), n.time_num - 1), hi = std::min(lo + 1, n.time_num - 1);
            const float frac = (lo == hi) ? 0.f : t - std::floor(t);
            P.time_value = t;
            const DeviceGrid& A = h->d_time[lo]; const DeviceGrid& B = h->d_time[hi];
            P.gx = A.X; P.gy = A.Y; P.gz = A.Z;
            const bool exact = h->kernel_mode == FVSRN_KERNEL_REFERENCE_EXACT && !A.tex32.empty() && !B.tex32.empty();
            for (int i = 0; i < A.C / 16; ++i, ++g) {
                P.groups[g] = GridGroup{grid_ptr(A, i), grid_ptr(B, i), frac, A.encoding ? A.C / 16 : A.C * 2 / 16};
                fill_encoding(P, g, A, B, i);
                for (int q = 0; q < 4; ++q) {
                    P.tex_a[4 * g + q] = exact ? A.tex32[4 * i + q] : A.tex[4 * i + q];
                    P.tex_b[4 * g + q] = exact ? B.tex32[4 * i + q] : B.tex[4 * i + q];
                }
            }
        }
        if (!h->d_ens.empty()) {
            const DeviceGrid& A = h->d_ens[n.interpolate_ensemble(ensemble)];
            P.gx = A.X; P.gy = A.Y; P.gz = A.Z;
            const bool exact = h->kernel_mode == FVSRN_KERNEL_REFERENCE_EXACT && !A.tex32.empty();
            for (int i = 0; i < A.C / 16; ++i, ++g) {
                P.groups[g] = GridGroup{grid_ptr(A, i), grid_ptr(A, i), 0.f, A.encoding ? A.C / 16 : A.C * 2 / 16};
                fill_encoding(P, g, A, A, i);
                for (int q = 0; q < 4; ++q) P.tex_a[4 * g + q] = P.tex_b[4 * g + q] = exact ? A.tex32[4 * i + q] : A.tex[4 * i + q];
            }
        }
    }
    for (int i = 0; i < 3; ++i) { P.box_min[i] = unit_box ? 0.f : n.box_min[i]; P.box_size[i] = unit_box ? 1.f : n.box_size[i]; }
    if (n.fourier.size() > 2 * sizeof(P.fourier2) / sizeof(P.fourier2[0])) throw std::runtime_error("Fourier matrix too large");
    std::memcpy(P.fourier2, n.fourier.data(), n.fourier.size() * 2);
    {   // NeRF-style Fourier matrix? B[i][i%3] = 2^(i/3) * B[i%3][i%3] (exactly, in fp16), all other entries zero
        const int nf = n.num_fourier;
        bool nerf = !n.direction_in_fourier && nf >= 3 && int(n.fourier.size()) == 3 * nf;
        for (int i = 0; nerf && i < nf; ++i)
            for (int a = 0; a < 3; ++a) {
                const float v = f16_to_f32(n.fourier[i + size_t(nf) * a]);
                const float want = (a == i % 3) ? std::ldexp(f16_to_f32(n.fourier[(i % 3) + size_t(nf) * (i % 3)]), i / 3) : 0.f;
                if (v != want || (a == i % 3 && !(std::fabs(v) > 0.f && std::fabs(v) < 30000.f))) nerf = false;
            }
        P.fourier_nerf = nerf ? 1 : 0;
    }
    {   // last layer of density networks (cout == 1) for the in-register dot product of the tcgen05 kernels
        const Layer& last = n.layers.back();
        uint16_t* w = reinterpret_cast<uint16_t*>(P.wlast2);
        for (int k = 0; k < last.cin && k < 128; ++k) w[k] = last.w(0, k);
        P.blast = f16_to_f32(last.bias[0]);
    }
    return P;
}

void require_prepared(fvsrn_network* h) {
    if (!h) throw std::runtime_error("null network handle");
    if (!h->prepared) throw std::runtime_error("network is not prepared: call fvsrn_network_prepare first");
}

unsigned int* next_tile_counter(fvsrn_network* h, cudaStream_t s) {
    unsigned int* c = h->d_tile_counter + h->tile_counter_next;
    h->tile_counter_next = (h->tile_counter_next + 1) % h->tile_counter_slots;
    CUDA_OK(cudaMemsetAsync(c, 0, sizeof(unsigned int), s));
    return c;
}

} // namespace

// shared with api_volume.cu
namespace fvsrn_api {
int fail(const std::string& msg) { return ::fail(msg); }
DvrParams fill_dvr_params(const fvsrn_dvr_params* p, float* out, bool is_rgbo, const void* any_dev);
}

// =========================================================================================== C ABI
extern "C" {

int fvsrn_abi_version(void) { return FVSRN_ABI_VERSION; }
const char* fvsrn_last_error(void) { return g_last_error.c_str(); }

int fvsrn_network_create(fvsrn_network** out) {
    FV_TRY
    if (!out) return fail("out is null");
    *out = new fvsrn_network();
    return 0;
    FV_CATCH
}

void fvsrn_network_free(fvsrn_network* net) {
    if (!net) return;
    release_gpu(net);
    delete net;
}

int fvsrn_network_load_volnet(const void* bytes, size_t n, fvsrn_network** out) {
    FV_TRY
    if (!bytes || !out) return fail("null argument");
    fvsrn_network* h = new fvsrn_network();
    try { h->net = Network::load(static_cast<const uint8_t*>(bytes), n); }
    catch (...) { delete h; throw; }
    *out = h;
    return 0;
    FV_CATCH
}

int fvsrn_network_save_volnet(const fvsrn_network* net, void* buf, size_t cap, size_t* written) {
    FV_TRY
    if (!net || !written) return fail("null argument");
    const std::vector<uint8_t> b = net->net.save();
    *written = b.size();
    if (!buf) return 0;
    if (cap < b.size()) return fail("buffer too small");
    std::memcpy(buf, b.data(), b.size());
    return 0;
    FV_CATCH
}

int fvsrn_network_set_input(fvsrn_network* net, int has_time, int has_direction) {
    FV_TRY
    if (!net) return fail("null network handle");
    net->net.has_time = has_time != 0; net->net.has_direction = has_direction != 0;
    return 0;
    FV_CATCH
}

int fvsrn_network_set_fourier(fvsrn_network* net, const float* B, int nf, int cin, int premultiplied) {
    FV_TRY
    if (!net || !B) return fail("null argument");
    Network& n = net->net;
    if (cin == 3) n.direction_in_fourier = false;
    else if (cin == 6) {
        if (!n.has_direction) return fail("hasDirection==false, but the fourier matrix has input channels for the direction");
        n.direction_in_fourier = true;
    } else return fail("Unrecognized number of input channels. Actual: " + std::to_string(cin) + ", expected: 3 or 6");
    n.num_fourier = nf;
    n.fourier.assign(size_t(nf) * cin, 0);
    for (int o = 0; o < nf; ++o) for (int i = 0; i < cin; ++i) // reference :150-154 (double 2*pi times float, then to half)
        n.fourier[o + size_t(nf) * i] = f32_to_f16(float((premultiplied ? 1.0 : 2.0 * M_PI) * double(B[size_t(o) * cin + i])));
    return 0;
    FV_CATCH
}

int fvsrn_network_disable_fourier(fvsrn_network* net) {
    if (!net) return fail("null network handle");
    net->net.num_fourier = 0; net->net.direction_in_fourier = false; net->net.fourier.clear();
    return 0;
}

int fvsrn_network_set_output_mode(fvsrn_network* net, const char* mode) {
    FV_TRY
    if (!net || !mode) return fail("null argument");
    net->net.output_mode = output_mode_from_string(mode);
    return 0;
    FV_CATCH
}

int fvsrn_network_add_layer(fvsrn_network* net, const float* W, const float* b, int cout, int cin,
                            const char* activation, float activation_param) {
    FV_TRY
    if (!net || !W || !b || !activation) return fail("null argument");
    if (cout <= 0 || cin <= 0) return fail("bad layer size");
    Layer l; l.cin = cin; l.cout = cout;
    l.activation = activation_from_string(activation); l.activation_param = activation_param;
    l.weights.resize(size_t(cin) * cout); l.bias.resize(cout);
    for (size_t i = 0; i < l.weights.size(); ++i) l.weights[i] = f32_to_f16(W[i]);
    for (int i = 0; i < cout; ++i) l.bias[i] = f32_to_f16(b[i]);
    net->net.add_layer(l);
    return 0;
    FV_CATCH
}

int fvsrn_network_set_latent_grid_layout(fvsrn_network* net, int time_min, int time_num, int time_step,
                                         int ensemble_min, int ensemble_num) {
    FV_TRY
    if (!net) return fail("null network handle");
    if (time_num < 0 || ensemble_num < 0 || time_step <= 0) return fail("bad latent grid layout");
    Network& n = net->net;
    n.has_latent_grid = true;
    n.time_min = time_min; n.time_num = time_num; n.time_step = time_step;
    n.ensemble_min = ensemble_min; n.ensemble_num = ensemble_num;
    n.time_grids.assign(time_num, LatentGrid()); n.ensemble_grids.assign(ensemble_num, LatentGrid());
    return 0;
    FV_CATCH
}

int fvsrn_network_set_time_grid(fvsrn_network* net, int index, const float* grid, int C, int Z, int Y, int X,
                                int encoding, double* encoding_error) {
    FV_TRY
    if (!net || !grid) return fail("null argument");
    if (index < 0 || index >= int(net->net.time_grids.size())) return fail("index out of bounds!");
    const double e = net->net.time_grids[index].from_dense(grid, C, Z, Y, X, encoding);
    if (encoding_error) *encoding_error = e;
    return 0;
    FV_CATCH
}

int fvsrn_network_set_ensemble_grid(fvsrn_network* net, int index, const float* grid, int C, int Z, int Y, int X,
                                    int encoding, double* encoding_error) {
    FV_TRY
    if (!net || !grid) return fail("null argument");
    if (index < 0 || index >= int(net->net.ensemble_grids.size())) return fail("index out of bounds!");
    const double e = net->net.ensemble_grids[index].from_dense(grid, C, Z, Y, X, encoding);
    if (encoding_error) *encoding_error = e;
    return 0;
    FV_CATCH
}

int fvsrn_network_set_box(fvsrn_network* net, const float box_min[3], const float box_size[3]) {
    if (!net || !box_min || !box_size) return fail("null argument");
    for (int i = 0; i < 3; ++i) { net->net.box_min[i] = box_min[i]; net->net.box_size[i] = box_size[i]; }
    return 0;
}
int fvsrn_network_get_box(const fvsrn_network* net, float box_min[3], float box_size[3]) {
    if (!net || !box_min || !box_size) return fail("null argument");
    for (int i = 0; i < 3; ++i) { box_min[i] = net->net.box_min[i]; box_size[i] = net->net.box_size[i]; }
    return 0;
}
int fvsrn_network_valid(const fvsrn_network* net) {
    if (!net) return 0;
    std::string why;
    const bool ok = net->net.valid(&why);
    if (!ok) g_last_error = why;
    return ok ? 1 : 0;
}
int fvsrn_network_num_parameters(const fvsrn_network* net) { return net ? net->net.num_parameters() : 0; }
int fvsrn_network_compute_max_warps(const fvsrn_network* net, int only_shared, int adjoint) {
    return net ? net->net.compute_max_warps(only_shared != 0, adjoint != 0) : -1;
}
int fvsrn_network_set_time_and_ensemble(fvsrn_network* net, float time, int ensemble) {
    if (!net) return fail("null network handle");
    net->net.set_time_and_ensemble(time, ensemble);
    return 0;
}

int fvsrn_network_set_gradient_mode(fvsrn_network* net, int gradient_mode, float finite_differences_stepsize) {
    if (!net) return fail("null network handle");
    if (gradient_mode != FVSRN_GRADIENT_OFF_OR_DIRECT && gradient_mode != FVSRN_GRADIENT_FINITE_DIFFERENCES && gradient_mode != FVSRN_GRADIENT_ADJOINT_METHOD)
        return fail("unknown gradient mode");
    if (gradient_mode == FVSRN_GRADIENT_FINITE_DIFFERENCES && !(finite_differences_stepsize > 0.f)) return fail("finite_differences_stepsize must be positive");
    std::lock_guard<std::mutex> lock(net->mu);
    net->gradient_mode = gradient_mode;
    if (gradient_mode == FVSRN_GRADIENT_ADJOINT_METHOD) net->adjoint_grid_step = finite_differences_stepsize > 0.f ? finite_differences_stepsize : 0.f;
    else if (finite_differences_stepsize > 0.f) net->fd_stepsize = finite_differences_stepsize;
    return 0;
}

int fvsrn_network_set_kernel_mode(fvsrn_network* net, int mode) {
    if (!net) return fail("null network handle");
    if (mode != FVSRN_KERNEL_MMA_SYNC && mode != FVSRN_KERNEL_TCGEN05 && mode != FVSRN_KERNEL_REFERENCE_EXACT) return fail("unknown kernel mode");
    net->kernel_mode = mode;
    return 0;
}
int fvsrn_network_get_kernel_mode(const fvsrn_network* net) { return net ? net->kernel_mode : -1; }

int fvsrn_network_get_info(const fvsrn_network* net, fvsrn_network_info* info) {
    if (!net || !info) return fail("null argument");
    const Network& n = net->net;
    std::memset(info, 0, sizeof(*info));
    info->has_time = n.has_time; info->has_direction = n.has_direction; info->num_fourier = n.num_fourier;
    info->direction_in_fourier = n.direction_in_fourier; info->output_mode = n.output_mode;
    info->num_layers = int(n.layers.size());
    info->hidden_channels = n.layers.empty() ? 0 : n.layers[0].cout;
    info->latent_channels = n.latent_channels();
    const LatentGrid* g = !n.time_grids.empty() ? &n.time_grids[0] : (!n.ensemble_grids.empty() ? &n.ensemble_grids[0] : nullptr);
    if (g) { info->grid_x = g->X; info->grid_y = g->Y; info->grid_z = g->Z; info->grid_encoding = g->encoding; }
    info->time_min = n.time_min; info->time_num = n.time_num; info->time_step = n.time_step;
    info->ensemble_min = n.ensemble_min; info->ensemble_num = n.ensemble_num;
    info->activation = n.layers.empty() ? ACT_NONE : n.layers[0].activation;
    info->activation_param = n.layers.empty() ? 1.f : n.layers[0].activation_param;
    info->channels_out = kOutputModeChannelsOut[n.output_mode];
    info->gradient_mode = net->gradient_mode; info->finite_differences_stepsize = net->fd_stepsize;
    return 0;
}

int fvsrn_network_prepare(fvsrn_network* net, int device, void* stream) {
    FV_TRY
    if (!net) return fail("null network handle");
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0)
        return fail("no CUDA device available: the fV-SRN B200 path has no CPU fallback");
    if (device < 0 || device >= count) return fail("invalid device index");
    std::lock_guard<std::mutex> lock(net->mu);
    release_gpu(net);
    std::vector<uint8_t> blob;
    build_blob(net, blob);
    DeviceGuard guard(device);
    net->device = device;
    cudaDeviceProp prop;
    CUDA_OK(cudaGetDeviceProperties(&prop, device));
    if (prop.major < 10) return fail("the fV-SRN B200 kernels are built for sm_100a only; found sm_" + std::to_string(prop.major * 10 + prop.minor));
    net->sm_count = prop.multiProcessorCount;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    CUDA_OK(cudaMalloc(&net->d_blob, blob.size()));
    CUDA_OK(cudaMemcpyAsync(net->d_blob, blob.data(), blob.size(), cudaMemcpyHostToDevice, s));
    std::vector<uint8_t> tc_blob;
    build_tc_blob(net, tc_blob);
    CUDA_OK(cudaMalloc(&net->d_tc_blob, tc_blob.size()));
    CUDA_OK(cudaMemcpyAsync(net->d_tc_blob, tc_blob.data(), tc_blob.size(), cudaMemcpyHostToDevice, s));
    CUDA_OK(cudaStreamSynchronize(s)); // blobs are locals
    auto make_texture = [&](DeviceGrid& d, const void* host, const cudaChannelFormatDesc& fd, size_t texel_bytes, bool normalized_float, int X, int Y, int Z,
                            std::vector<cudaArray_t>& arrays, std::vector<cudaTextureObject_t>& texs) {
        cudaArray_t arr = nullptr;
        CUDA_OK(cudaMalloc3DArray(&arr, &fd, make_cudaExtent(X, Y, Z)));
        arrays.push_back(arr);
        cudaMemcpy3DParms cp = {};
        cp.srcPtr = make_cudaPitchedPtr(const_cast<void*>(host), size_t(X) * texel_bytes, X, Y);
        cp.dstArray = arr; cp.extent = make_cudaExtent(X, Y, Z); cp.kind = cudaMemcpyHostToDevice;
        CUDA_OK(cudaMemcpy3D(&cp));
        cudaResourceDesc rd = {}; rd.resType = cudaResourceTypeArray; rd.res.array.array = arr;
        cudaTextureDesc td = {};
        td.addressMode[0] = td.addressMode[1] = td.addressMode[2] = cudaAddressModeClamp;
        td.filterMode = cudaFilterModeLinear; td.readMode = normalized_float ? cudaReadModeNormalizedFloat : cudaReadModeElementType; td.normalizedCoords = 1;
        cudaTextureObject_t t = 0;
        CUDA_OK(cudaCreateTextureObject(&t, &rd, &td, nullptr));
        texs.push_back(t);
    };
    auto upload = [&](const LatentGrid& g) {
        DeviceGrid d; d.C = g.C; d.Z = g.Z; d.Y = g.Y; d.X = g.X; d.encoding = g.encoding;
        const size_t voxels = size_t(g.Z) * g.Y * g.X;
        if (g.encoding != GRID_FLOAT) {
            // Byte-encoded grid: it stays 8-bit on the device. C/4 unorm8 RGBA textures read as normalised floats - the
            // reference's own layout (LatentGrid::GPUArray with isFloat = false, :470-504) - and a channel-last byte copy
            // for the mma.sync kernels; decoded after the filter (srn_core.cuh decode_grid_value).
            std::vector<uint8_t> cl(voxels * g.C);
            for (int z = 0; z < g.Z; ++z) for (int y = 0; y < g.Y; ++y) for (int x = 0; x < g.X; ++x)
                for (int c = 0; c < g.C; ++c)
                    cl[((size_t(z) * g.Y + y) * g.X + x) * g.C + c] = uint8_t(g.data[g.idx(c / 4, z, y, x, c % 4)]);
            CUDA_OK(cudaMalloc(&d.data, cl.size()));
            CUDA_OK(cudaMemcpyAsync(d.data, cl.data(), cl.size(), cudaMemcpyHostToDevice, s));
            CUDA_OK(cudaStreamSynchronize(s));
            const cudaChannelFormatDesc fd = cudaCreateChannelDesc(8, 8, 8, 8, cudaChannelFormatKindUnsigned);
            for (int q = 0; q < g.C / 4; ++q)       // storage order of the .volnet grid is already (C/4, Z, Y, X, 4)
                make_texture(d, g.data.data() + size_t(q) * voxels * 4, fd, 4, true, g.X, g.Y, g.Z, d.arrays, d.tex);
            d.offset = g.offset_or_mean; d.scale = g.scale_or_std;
            return d;
        }
        const std::vector<uint16_t> cl = grid_channel_last(g);
        CUDA_OK(cudaMalloc(&d.data, cl.size() * 2));
        CUDA_OK(cudaMemcpyAsync(d.data, cl.data(), cl.size() * 2, cudaMemcpyHostToDevice, s));
        CUDA_OK(cudaStreamSynchronize(s)); // cl is a temporary
        // the same voxels as C/4 half4 textures: clamp addressing, linear filter, normalised coordinates
        // (reference LatentGrid::GPUArray, volume_interpolation_network.cpp:470-504, with fp16 instead of fp32 texels)
        std::vector<uint16_t> sub(voxels * 4);
        for (int q = 0; q < g.C / 4; ++q) {
            for (size_t v = 0; v < voxels; ++v)
                for (int c = 0; c < 4; ++c) sub[v * 4 + c] = cl[v * g.C + q * 4 + c];
            make_texture(d, sub.data(), cudaCreateChannelDescHalf4(), 8, false, g.X, g.Y, g.Z, d.arrays, d.tex);
        }
        return d;
    };
    if (net->net.has_latent_grid) {
        for (const auto& g : net->net.time_grids) net->d_time.push_back(upload(g));
        for (const auto& g : net->net.ensemble_grids) net->d_ens.push_back(upload(g));
        const DeviceGrid* r = !net->d_time.empty() ? &net->d_time[0] : &net->d_ens[0];
        for (const auto& g : net->d_time) if (g.X != r->X || g.Y != r->Y || g.Z != r->Z) return fail("all latent grids must share one resolution");
        for (const auto& g : net->d_ens) if (g.X != r->X || g.Y != r->Y || g.Z != r->Z) return fail("all latent grids must share one resolution");
    }
    if (net->d_time.size() > 1 && net->d_time[0].encoding == GRID_FLOAT) {
        // scratch grids for blended keyframe pairs (4: consecutive views alternate two streams, so two are in flight)
        const DeviceGrid& r = net->d_time[0];
        for (int k = 0; k < 4; ++k) {
            DeviceGrid d; d.C = r.C; d.Z = r.Z; d.Y = r.Y; d.X = r.X; d.encoding = GRID_FLOAT;
            CUDA_OK(cudaMalloc(&d.data, size_t(r.C) * r.Z * r.Y * r.X * 2));
            const cudaChannelFormatDesc fd = cudaCreateChannelDescHalf4();
            for (int q = 0; q < r.C / 4; ++q) {
                cudaArray_t arr = nullptr;
                CUDA_OK(cudaMalloc3DArray(&arr, &fd, make_cudaExtent(r.X, r.Y, r.Z), cudaArraySurfaceLoadStore));
                d.arrays.push_back(arr);
                cudaResourceDesc rd = {}; rd.resType = cudaResourceTypeArray; rd.res.array.array = arr;
                cudaTextureDesc td = {};
                td.addressMode[0] = td.addressMode[1] = td.addressMode[2] = cudaAddressModeClamp;
                td.filterMode = cudaFilterModeLinear; td.readMode = cudaReadModeElementType; td.normalizedCoords = 1;
                cudaTextureObject_t t = 0; cudaSurfaceObject_t sf = 0;
                CUDA_OK(cudaCreateTextureObject(&t, &rd, &td, nullptr));
                CUDA_OK(cudaCreateSurfaceObject(&sf, &rd));
                d.tex.push_back(t); d.surf.push_back(sf);
            }
            for (auto& e : d.ev_read) CUDA_OK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
            net->d_blend.push_back(std::move(d));
        }
    }
    net->tile_counter_slots = 256;
    CUDA_OK(cudaMalloc(&net->d_tile_counter, sizeof(unsigned int) * net->tile_counter_slots));
    CUDA_OK(cudaMemsetAsync(net->d_tile_counter, 0, sizeof(unsigned int) * net->tile_counter_slots, s));
    CUDA_OK(cudaStreamSynchronize(s));
    CUDA_OK(cudaStreamCreateWithFlags(&net->aux_stream, cudaStreamNonBlocking));
    CUDA_OK(cudaEventCreateWithFlags(&net->ev_fork, cudaEventDisableTiming));
    CUDA_OK(cudaEventCreateWithFlags(&net->ev_join, cudaEventDisableTiming));
    net->prepared = true;
    return 0;
    FV_CATCH
}

int fvsrn_network_clear_gpu_resources(fvsrn_network* net) {
    if (!net) return fail("null network handle");
    std::lock_guard<std::mutex> lock(net->mu);
    release_gpu(net);
    return 0;
}

} // extern "C"

// =========================================================================================== compute entry points
namespace {

__global__ void extract_color_kernel(const float* __restrict__ img, float* __restrict__ out, long long hw, long long total) {
    // (B,8,H,W) -> (B,4,H,W): reference SelectOutputChannelKernel, ChannelColor branch (iimage_evaluator_cuda.cu:21-80)
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const long long b = i / (4 * hw), r = i - b * 4 * hw;
        out[i] = img[b * 8 * hw + r];
    }
}

// IImageEvaluator::ExtractColor on a raw (B,8,H,W) tensor: SelectOutputChannelKernel2 / TonemappingKernel2 and their
// packed-texture forms (iimage_evaluator_cuda.cu:21-80, 168-260). mm = {min depth, max depth} for ChannelDepth.
__global__ void extract_color_ex_kernel(const float* __restrict__ img, void* __restrict__ out, long long hw, int batch,
                                        int channel_mode, int tonemap, float max_exposure, int fmt, const float* __restrict__ mm) {
    const long long total = hw * batch;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const long long b = i / hw, px = i - b * hw;
        const float* p = img + b * 8 * hw + px;
        const float4 color = make_float4(p[0], p[hw], p[2 * hw], p[3 * hw]);
        float4 c;
        if (channel_mode == FVSRN_CHANNEL_DEPTH) {      // iimage_evaluator.cpp:84-94
            const float mn = mm[0], mx = mm[1];
            const float scale = 1.f / (mx - mn), offset = -mn / (mx - mn);
            const float d = p[7 * hw] * scale + offset;
            c = make_float4(d, d, d, color.w * 0.f + 1.f);
        } else {
            c = select_channel(channel_mode, tonemap, max_exposure, color, make_float3(p[4 * hw], p[5 * hw], p[6 * hw]));
        }
        if (fmt == FVSRN_OUTPUT_RGBA_F32) {
            float* o = reinterpret_cast<float*>(out) + b * 4 * hw + px;
            o[0] = c.x; o[hw] = c.y; o[2 * hw] = c.z; o[3 * hw] = c.w;
        } else if (fmt == FVSRN_OUTPUT_RGBA8) {
            reinterpret_cast<unsigned int*>(out)[i] = rgba_to_int(c.x, c.y, c.z, c.w);
        } else {
            reinterpret_cast<uint2*>(out)[i] = make_uint2(h2u(__floats2half2_rn(c.x, c.y)), h2u(__floats2half2_rn(c.z, c.w)));
        }
    }
}
// min / max of the depth plane with torch.min / torch.max semantics (NaN propagates): {min, max} -> mm
__global__ void depth_minmax_kernel(const float* __restrict__ img, long long hw, int batch, float* mm) {
    __shared__ float smin[32], smax[32];
    __shared__ int snan[32];
    float mn = FLT_MAX, mx = -FLT_MAX; int nan = 0;
    const long long total = hw * batch;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const long long b = i / hw, px = i - b * hw;
        const float d = img[b * 8 * hw + 7 * hw + px];
        if (d != d) nan = 1; else { mn = fminf(mn, d); mx = fmaxf(mx, d); }
    }
    for (int o = 16; o > 0; o >>= 1) {
        mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, o)); mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        nan |= __shfl_xor_sync(0xffffffffu, nan, o);
    }
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    if (l == 0) { smin[w] = mn; smax[w] = mx; snan[w] = nan; }
    __syncthreads();
    if (w == 0) {
        const int nw = blockDim.x >> 5;
        mn = l < nw ? smin[l] : FLT_MAX; mx = l < nw ? smax[l] : -FLT_MAX; nan = l < nw ? snan[l] : 0;
        for (int o = 16; o > 0; o >>= 1) {
            mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, o)); mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
            nan |= __shfl_xor_sync(0xffffffffu, nan, o);
        }
        if (l == 0) {   // float atomics through the ordered-int trick are overkill here: one block per call (grid = 1)
            mm[0] = nan ? __int_as_float(0x7fc00000) : mn; mm[1] = nan ? __int_as_float(0x7fc00000) : mx;
        }
    }
}

void ensure_scratch(fvsrn_network* h, size_t bytes) {
    if (h->scratch_bytes >= bytes) return;
    if (h->d_scratch) CUDA_OK(cudaFree(h->d_scratch));
    h->d_scratch = nullptr; h->scratch_bytes = 0;
    CUDA_OK(cudaMalloc(&h->d_scratch, bytes));
    h->scratch_bytes = bytes;
}

// texels of a keyframe pair blended in fp32 in the reference's order (a + frac (b - a), renderer_volume_tensorcores.cuh:590-604),
// rounded to fp16 once; one thread per voxel and 4-channel quad: the channel-last copy and the quad's half4 3-D array
struct BlendSurfaces { unsigned long long s[8]; };
__global__ void blend_keyframes_kernel(const uint2* __restrict__ a, const uint2* __restrict__ b, float frac, uint2* __restrict__ out,
                                       const __grid_constant__ BlendSurfaces S, int quads, int X, int Y, long long total) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const uint2 va = a[i], vb = b[i];
        const float2 a0 = __half22float2(u2h(va.x)), a1 = __half22float2(u2h(va.y));
        const float2 b0 = __half22float2(u2h(vb.x)), b1 = __half22float2(u2h(vb.y));
        uint2 r;
        r.x = h2u(__floats2half2_rn(fmaf(frac, b0.x - a0.x, a0.x), fmaf(frac, b0.y - a0.y, a0.y)));
        r.y = h2u(__floats2half2_rn(fmaf(frac, b1.x - a1.x, a1.x), fmaf(frac, b1.y - a1.y, a1.y)));
        out[i] = r;
        const long long voxel = i / quads; const int q = int(i - voxel * quads);
        const int x = int(voxel % X), y = int((voxel / X) % Y), z = int(voxel / ((long long)X * Y));
        surf3Dwrite(r, (cudaSurfaceObject_t)S.s[q], x * 8, y, z);
    }
}
// tcgen05 kernels with time-interpolated grids: replace the keyframe pair of P by one scratch grid holding their blend, so
// that the kernel fetches one grid per sample (SPEC 1 / the LSU eval path) instead of two. The fp16 rounding of the blended
// texel is the only difference to blending after the filter (both are linear); measured on the C5 frame against the reference
// kernel in tests/test_gpu_baseline_configs.py. Returns the scratch grid (one of its ev_read must be recorded after the consuming
// kernel) or nullptr when nothing was replaced (static grid, byte grids, other kernel families).
DeviceGrid* blend_keyframes(fvsrn_network* h, NetParams& P, float time_abs, cudaStream_t s) {
    const Network& n = h->net;
    if (h->d_blend.empty() || h->d_time.size() < 2 || P.grid_encoding != 0) return nullptr;
    const float t = n.interpolate_time(time_abs);
    const int lo = std::min(int(t), n.time_num - 1), hi = std::min(lo + 1, n.time_num - 1);
    const float frac = (lo == hi) ? 0.f : t - std::floor(t);
    if (!(frac > 0.f)) return nullptr;
    DeviceGrid* d = nullptr;
    for (auto& c : h->d_blend) if (c.key_lo == lo && c.key_hi == hi && c.key_frac == frac) d = &c;
    if (!d) {
        d = &h->d_blend[size_t(h->blend_next)];
        h->blend_next = (h->blend_next + 1) % int(h->d_blend.size());
        for (auto e : d->ev_read) CUDA_OK(cudaStreamWaitEvent(s, e, 0));   // the last kernels that read this scratch grid, on either stream
        const DeviceGrid& A = h->d_time[size_t(lo)]; const DeviceGrid& B = h->d_time[size_t(hi)];
        BlendSurfaces S{};
        const int quads = A.C / 4;
        if (quads > 8) return nullptr;
        for (int q = 0; q < quads; ++q) S.s[q] = d->surf[size_t(q)];
        const long long total = (long long)A.X * A.Y * A.Z * quads;
        const int blocks = int(std::min<long long>((total + 255) / 256, 148 * 8));
        blend_keyframes_kernel<<<blocks, 256, 0, s>>>(static_cast<const uint2*>(A.data), static_cast<const uint2*>(B.data), frac,
                                                      static_cast<uint2*>(d->data), S, quads, A.X, A.Y, total);
        CUDA_OK(cudaGetLastError());
        h->launches++;
        d->key_lo = lo; d->key_hi = hi; d->key_frac = frac;
    }
    for (int i = 0; i < d->C / 16; ++i) {       // the time grids are the first groups of P (make_net_params)
        P.groups[i] = GridGroup{grid_ptr(*d, i), grid_ptr(*d, i), 0.f, d->C * 2 / 16};
        for (int q = 0; q < 4; ++q) P.tex_a[4 * i + q] = P.tex_b[4 * i + q] = d->tex[size_t(4 * i + q)];
    }
    return d;
}

void do_evaluate(fvsrn_network* h, const float* pos, const float* dir, float* out, int64_t n, cudaStream_t s) {
    require_prepared(h);
    if (n < 0) throw std::runtime_error("negative sample count");
    if (n == 0) return;
    if (!pos || !out) throw std::runtime_error("null device pointer");
    if (h->net.has_direction && !dir) throw std::runtime_error("the network uses directions, but no directions were passed");
    if (h->kernel_mode == FVSRN_KERNEL_REFERENCE_EXACT) throw std::runtime_error("the reference-exact kernel mode covers the DVR frame only");
    // reference volume_interpolation.cpp:45-48: positions are in the unit cube
    // the tcgen05 kernels cover the 1- and 4-channel outputs; gradient / curvature networks run on the mma.sync family
    const bool tc = h->kernel_mode == FVSRN_KERNEL_TCGEN05 && h->net.output_mode <= OUT_RGBO_DIRECT;
    NetParams P = make_net_params(h, h->net.current_time, h->net.current_ensemble, true, tc);
    DeviceGrid* blended = tc ? blend_keyframes(h, P, h->net.current_time, s) : nullptr;
    EvalParams E{pos, dir, out, (long long)n, kOutputModeChannelsOut[h->net.output_mode]};
    LaunchCfg cfg; cfg.sm_count = h->sm_count;
    cfg.max_blocks = tc ? (n + 2 * kTcComputeThreads - 1) / (2 * kTcComputeThreads) : ((n + 31) / 32 + kWarpsPerBlock - 1) / kWarpsPerBlock;
    launch_fn fn = pick_launcher(h->H16);
    if (!fn) throw std::runtime_error("no kernel for this hidden width");
    CUDA_OK(fn(tc ? KIND_EVAL_TC : KIND_EVAL, h->GC16, h->dir, P, &E, &cfg, s));
    if (blended) CUDA_OK(cudaEventRecord(blended->ev_read[0], s));
    h->launches++;
}

// fvsrn_dvr_params -> the kernels' launch parameters (validation shared by the SRN and the grid renderer);
// is_rgbo: colour networks skip the transfer function (SKIP_TRANSFER_FUNCTION), `any_dev` is some readable device pointer
DvrParams fill_dvr_params(const fvsrn_dvr_params* p, float* out, bool is_rgbo, const void* any_dev) {
    if (!p || !out) throw std::runtime_error("null argument");
    if (p->width <= 0 || p->height <= 0 || p->batch <= 0) throw std::runtime_error("image size and batch must be positive");
    if (!p->camera_matrix_dev) throw std::runtime_error("camera matrix is null");
    if (!(p->stepsize > 0.f)) throw std::runtime_error("stepsize must be positive");
    if (!is_rgbo) {
        if (!p->tf_dev) throw std::runtime_error("texture tensor is not defined");
        if (p->tf_kind < 0 || p->tf_kind > 2) throw std::runtime_error("unknown transfer function kind");
        if (p->tf_kind != FVSRN_TF_IDENTITY && p->tf_resolution < 2) throw std::runtime_error("transfer function needs at least 2 entries");
        const int floats = p->tf_kind == 0 ? p->tf_resolution * 5 : (p->tf_kind == 1 ? p->tf_resolution * 4 : 2);
        if (floats > kTfSmemFloats) throw std::runtime_error("transfer function too large for the fused kernel (max 204 control points / 256 texels)");
        if (p->tf_batch != 1 && p->tf_batch != p->batch) throw std::runtime_error("transfer function batch must be 1 or equal to the camera batch");
    }
    int spp = p->samples_per_pixel <= 0 ? 1 : p->samples_per_pixel;
    int side = int(std::lround(std::sqrt(double(spp))));
    if (side * side != spp) throw std::runtime_error("samples_per_pixel must be a square number (1, 4, 9, ...)");
    DvrParams D{};
    D.W = p->width; D.H = p->height;
    D.x0 = p->tile_x0; D.y0 = p->tile_y0;
    D.tw = p->tile_w > 0 ? p->tile_w : p->width - p->tile_x0;
    D.th = p->tile_h > 0 ? p->tile_h : p->height - p->tile_y0;
    if (D.x0 < 0 || D.y0 < 0 || D.tw <= 0 || D.th <= 0 || D.x0 + D.tw > D.W || D.y0 + D.th > D.H) throw std::runtime_error("tile rectangle outside the image");
    D.cam = p->camera_matrix_dev;
    D.fov_y = p->fov_y_radians;
    D.aspect = float(double(p->width) / double(p->height)); // camera->setAspectRatio, image_evaluator_simple.cpp:204
    D.stepsize = p->stepsize; D.alpha_early_out = p->alpha_early_out;
    D.dmin = p->density_min; D.dmax = p->density_max; D.inv_range = 1.0f / (p->density_max - p->density_min);
    D.early_out = p->early_out; D.blend_mode = p->blend_mode;
    D.tf_kind = is_rgbo ? 2 : p->tf_kind; D.tf_R = p->tf_resolution; D.tf_batch = is_rgbo ? 1 : p->tf_batch;
    D.tf = is_rgbo ? reinterpret_cast<const float*>(any_dev) : p->tf_dev; // rgbo networks skip the TF; any readable pointer
    D.spp = spp; D.spp_side = side;
    D.out = out; D.sb = p->out_stride_b; D.sc = p->out_stride_c; D.sy = p->out_stride_y; D.sx = p->out_stride_x;
    D.step_count = p->step_count_dev; D.counters = p->counters_dev;
    D.out_format = p->output_format; D.channel_mode = p->channel_mode; D.tonemap = p->tonemap; D.max_exposure = p->max_exposure;
    if (p->brdf_flags & ~(FVSRN_BRDF_MAGNITUDE_SCALING | FVSRN_BRDF_PHONG)) throw std::runtime_error("unknown BRDF flags");
    if (p->brdf_flags) {
        if (is_rgbo) throw std::runtime_error("BRDF shading needs normals: colour networks provide none");
        if (p->light_type != FVSRN_LIGHT_POINT && p->light_type != FVSRN_LIGHT_DIRECTIONAL) throw std::runtime_error("unknown light type");
        D.brdf_flags = p->brdf_flags; D.light_type = p->light_type;
        for (int i = 0; i < 3; ++i) D.light[i] = p->light[i];
        D.brdf_mag_scaling = p->brdf_magnitude_scaling; D.brdf_ambient = p->brdf_ambient; D.brdf_specular = p->brdf_specular;
        D.brdf_mag_center = p->brdf_magnitude_center; D.brdf_mag_radius = p->brdf_magnitude_radius; D.brdf_spec_exp = p->brdf_specular_exponent;
    }
    if (D.out_format < FVSRN_OUTPUT_RAW8_F32 || D.out_format > FVSRN_OUTPUT_RGBA_F16) throw std::runtime_error("unknown output format");
    if (D.out_format != FVSRN_OUTPUT_RAW8_F32) {
        if (D.channel_mode == FVSRN_CHANNEL_DEPTH) throw std::runtime_error("ChannelDepth needs the frame's min/max depth: render the raw tensor and call fvsrn_extract_color_ex");
        if (D.channel_mode < 0 || D.channel_mode > FVSRN_CHANNEL_COLOR) throw std::runtime_error("unknown channel mode");
        if (D.tonemap && !(D.max_exposure > 0.f)) throw std::runtime_error("tonemapping needs a positive max_exposure");
    }
    return D;
}

} // namespace
namespace fvsrn_api {
DvrParams fill_dvr_params(const fvsrn_dvr_params* p, float* out, bool is_rgbo, const void* any_dev) { return ::fill_dvr_params(p, out, is_rgbo, any_dev); }
}
namespace {

void do_render(fvsrn_network* h, const fvsrn_dvr_params* p, float* out, cudaStream_t s) {
    require_prepared(h);
    const int om = h->net.output_mode;
    const bool is_rgbo = om == OUT_RGBO || om == OUT_RGBO_DIRECT;
    DvrParams D = fill_dvr_params(p, out, is_rgbo, h->d_blob);
    // normals: predicted by the network, finite differences, or none (the normal channels stay zero). The
    // reference runs its finite differences whenever the module asks for them, shaded or not
    // (volume_interpolation_network.cpp:1146 is commented out) - so do we.
    if (h->gradient_mode != FVSRN_GRADIENT_OFF_OR_DIRECT && is_rgbo) throw std::runtime_error("gradient modes need a density network (the reference's evalNormal asserts it)");
    D.normals = (om >= OUT_DENSITYGRAD || h->gradient_mode != FVSRN_GRADIENT_OFF_OR_DIRECT || D.brdf_flags) ? 1 : 0;
    // the tcgen05 kernels march density / colour networks without normals; everything with normals runs on the mma.sync family
    // Adjoint normals (GRADIENT_MODE 2): one backward pass through the network per shaded step - in the fast kernel modes on the
    // mma.sync family (srn_core.cuh Srn::adjoint), in reference-exact mode bit-identical to the reference (kernels_exact.cuh SrnExact::adjoint)
    const bool adjoint = h->gradient_mode == FVSRN_GRADIENT_ADJOINT_METHOD;
    if (adjoint) {
        const Network& n = h->net;
        if (n.has_direction) throw std::runtime_error("adjoint gradient mode: networks with a direction input are not supported (the reference's adjoint ignores the direction, :1478-1504)");
        if (om == OUT_DENSITYCURV || om == OUT_DENSITYCURV_DIRECT) throw std::runtime_error("adjoint gradient mode: not defined for densitycurvature networks (reference :1262-1264)");
        if (h->num_hidden > 7 && h->kernel_mode == FVSRN_KERNEL_REFERENCE_EXACT) throw std::runtime_error("adjoint gradient mode, reference-exact kernels: at most 7 hidden layers after the first");
        if (h->GC16 < 1) throw std::runtime_error("adjoint gradient mode: networks with a latent grid");
        for (const auto& g : h->d_time) if (g.encoding != GRID_FLOAT) throw std::runtime_error("adjoint gradient mode: float latent grids only");
        for (const auto& g : h->d_ens) if (g.encoding != GRID_FLOAT) throw std::runtime_error("adjoint gradient mode: float latent grids only");
    }
    const bool tc = h->kernel_mode == FVSRN_KERNEL_TCGEN05 && !D.normals;
    const bool exact = h->kernel_mode == FVSRN_KERNEL_REFERENCE_EXACT;
    if (exact) {    // the shapes the reference kernels this mode is validated against cover (kernels_exact.cuh)
        const Network& n = h->net;
        if (is_rgbo || h->GC16 < 1 || h->GC16 > 2 || ((n.has_direction || h->GC16 == 2) && h->H16 != 2 && h->H16 != 4))
            throw std::runtime_error("reference-exact mode: density-type networks with a latent grid; direction input or 32 latent channels with hidden width 32 or 64");
        for (size_t i = 0; i < h->d_time.size(); ++i) ensure_tex32(h->d_time[i], n.time_grids[i]);
        for (size_t i = 0; i < h->d_ens.size(); ++i) ensure_tex32(h->d_ens[i], n.ensemble_grids[i]);
    }
    launch_fn fn = pick_launcher(h->H16);
    if (!fn) throw std::runtime_error("no kernel for this hidden width");
    auto launch = [&](int b0, int nb, float time_abs, cudaStream_t ls) {
        NetParams P = make_net_params(h, time_abs, h->net.current_ensemble, false, tc);
        DeviceGrid* blended = tc ? blend_keyframes(h, P, time_abs, ls) : nullptr;
        DvrParams d = D; d.b0 = b0; d.nb = nb;      // the launcher lays the tile raster of its kernel over the rectangle
        d.tile_counter = next_tile_counter(h, ls);
        LaunchCfg cfg; cfg.sm_count = h->sm_count;
        if (adjoint && exact) CUDA_OK(launch_exact_adjoint(h->H16, h->GC16, P, d, &cfg, ls));
        else if (exact) CUDA_OK(launch_exact(h->H16, h->GC16, h->dir, P, d, &cfg, ls));
        else CUDA_OK(fn(tc ? KIND_DVR_TC : KIND_DVR, h->GC16, h->dir, P, &d, &cfg, ls));
        if (blended) CUDA_OK(cudaEventRecord(blended->ev_read[ls == h->aux_stream ? 1 : 0], ls));
        h->launches++;
    };
    if (p->view_times_host) {
        // One launch per view: every view selects its own keyframe pair (time sweeps). The persistent grid of
        // a view drains over its last tiles; the views alternate between the caller's stream and an internal one so that
        // the next view's CTAs become resident on the SMs the previous view is leaving (views write disjoint outputs).
        const bool overlap = p->batch > 1 && h->aux_stream != nullptr;
        if (overlap) { CUDA_OK(cudaEventRecord(h->ev_fork, s)); CUDA_OK(cudaStreamWaitEvent(h->aux_stream, h->ev_fork, 0)); }
        for (int b = 0; b < p->batch; ++b) {
            const float t = std::min(std::max(p->view_times_host[b], float(h->net.time_min)), float(h->net.time_max_inclusive()));
            launch(b, 1, t, (overlap && (b & 1)) ? h->aux_stream : s);
        }
        if (overlap) { CUDA_OK(cudaEventRecord(h->ev_join, h->aux_stream)); CUDA_OK(cudaStreamWaitEvent(s, h->ev_join, 0)); }
    } else launch(0, p->batch, h->net.current_time, s);
}

} // namespace

extern "C" {

int fvsrn_evaluate(fvsrn_network* net, const float* positions_dev, const float* directions_dev, float* out_dev,
                   int64_t n, void* stream) {
    FV_TRY
    if (!net) return fail("null network handle");
    std::lock_guard<std::mutex> lock(net->mu);
    DeviceGuard guard(net->device);
    do_evaluate(net, positions_dev, directions_dev, out_dev, n, static_cast<cudaStream_t>(stream));
    return 0;
    FV_CATCH
}

int fvsrn_evaluate_host(fvsrn_network* net, const float* positions_host, const float* directions_host, float* out_host,
                        int64_t n, void* stream) {
    FV_TRY
    if (!net) return fail("null network handle");
    std::lock_guard<std::mutex> lock(net->mu);
    DeviceGuard guard(net->device);
    require_prepared(net);
    if (n <= 0) return n == 0 ? 0 : fail("negative sample count");
    if (!positions_host || !out_host) return fail("null host pointer");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const int cout = kOutputModeChannelsOut[net->net.output_mode];
    const size_t pos_b = size_t(n) * 12, dir_b = directions_host ? pos_b : 0, out_b = size_t(n) * 4 * cout;
    ensure_scratch(net, pos_b + dir_b + out_b + 64);
    char* base = reinterpret_cast<char*>(net->d_scratch);
    float* d_pos = reinterpret_cast<float*>(base);
    float* d_dir = directions_host ? reinterpret_cast<float*>(base + pos_b) : nullptr;
    float* d_out = reinterpret_cast<float*>(base + ((pos_b + dir_b + 15) / 16) * 16);
    CUDA_OK(cudaMemcpyAsync(d_pos, positions_host, pos_b, cudaMemcpyHostToDevice, s));
    if (d_dir) CUDA_OK(cudaMemcpyAsync(d_dir, directions_host, dir_b, cudaMemcpyHostToDevice, s));
    do_evaluate(net, d_pos, d_dir, d_out, n, s);
    CUDA_OK(cudaMemcpyAsync(out_host, d_out, out_b, cudaMemcpyDeviceToHost, s));
    CUDA_OK(cudaStreamSynchronize(s));
    return 0;
    FV_CATCH
}

int fvsrn_render_dvr(fvsrn_network* net, const fvsrn_dvr_params* p, float* out_dev, void* stream) {
    FV_TRY
    if (!net) return fail("null network handle");
    std::lock_guard<std::mutex> lock(net->mu);
    DeviceGuard guard(net->device);
    do_render(net, p, out_dev, static_cast<cudaStream_t>(stream));
    return 0;
    FV_CATCH
}

int fvsrn_render_dvr_host(fvsrn_network* net, const fvsrn_dvr_params* p, float* out_host, void* stream) {
    FV_TRY
    if (!net) return fail("null network handle");
    std::lock_guard<std::mutex> lock(net->mu);
    DeviceGuard guard(net->device);
    require_prepared(net);
    if (!p || !out_host) return fail("null argument");
    // element size and planes of the output format (fvsrn_output_format): planar fp32 (8 or 4 planes) or packed pixels
    const int fmt = p->output_format;
    if (fmt < FVSRN_OUTPUT_RAW8_F32 || fmt > FVSRN_OUTPUT_RGBA_F16) return fail("unknown output format");
    const int planes = fmt == FVSRN_OUTPUT_RAW8_F32 ? 8 : (fmt == FVSRN_OUTPUT_RGBA_F32 ? 4 : 1);
    const size_t elem = fmt == FVSRN_OUTPUT_RGBA_F16 ? 8 : 4;
    const int64_t W = p->width, H = p->height;
    if (p->out_stride_x != 1 || p->out_stride_y != W || (planes > 1 && p->out_stride_c != W * H) || p->out_stride_b != planes * W * H)
        return fail("fvsrn_render_dvr_host needs a contiguous output: (B,8,H,W) / (B,4,H,W) fp32 or (B,H,W) packed pixels");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const bool is_rgbo = net->net.output_mode == OUT_RGBO || net->net.output_mode == OUT_RGBO_DIRECT;
    const size_t cam_b = size_t(p->batch) * 36;
    const int tf_floats = is_rgbo ? 0 : (p->tf_kind == 0 ? p->tf_resolution * 5 : (p->tf_kind == 1 ? p->tf_resolution * 4 : 2)) * p->tf_batch;
    const size_t tf_off = ((cam_b + 15) / 16) * 16, img_off = ((tf_off + size_t(tf_floats) * 4 + 255) / 256) * 256;
    const size_t img_b = size_t(p->batch) * planes * W * H * elem;
    ensure_scratch(net, img_off + img_b);
    char* base = reinterpret_cast<char*>(net->d_scratch);
    CUDA_OK(cudaMemcpyAsync(base, p->camera_matrix_dev, cam_b, cudaMemcpyHostToDevice, s));
    if (tf_floats) CUDA_OK(cudaMemcpyAsync(base + tf_off, p->tf_dev, size_t(tf_floats) * 4, cudaMemcpyHostToDevice, s));
    fvsrn_dvr_params q = *p;
    q.camera_matrix_dev = reinterpret_cast<const float*>(base);
    q.tf_dev = reinterpret_cast<const float*>(base + tf_off);
    char* d_img = base + img_off;
    const bool has_rect = q.tile_w > 0 || q.tile_h > 0 || q.tile_x0 > 0 || q.tile_y0 > 0;
    // A single full frame is rendered as a few row bands that alternate between the caller's stream and the internal one:
    // each band's device-to-host copy follows its kernel on the same stream and overlaps the next band's kernel on the
    // other, and the next band's CTAs fill the SMs the previous band is draining.
    // Pinned, device-mapped host memory (cudaHostAlloc / cudaHostRegister; every pinned allocation under unified addressing): the
    // kernel writes its pixels straight into the caller's buffer. The 4 - 32 bytes per pixel trickle over PCIe while the frame is
    // marched (16 MiB in 4.7 ms = 3.5 GB/s on C3) - no staging image, no copy left after the kernel.
    // A/B on one B200 (C3, kernel 4.67 ms; profiles/r02k_e2e_direct.log, r02q): the packed formats (4 / 8 bytes per pixel) 4.70 ms direct on
    // every box; fp32 colour (16 B/pixel in four planes) 4.7x ms direct on one box and 6.0 ms on another (4-byte posted writes are at
    // the mercy of the host's PCIe / NUMA path), 4.84 ms staged on both - so only the packed formats go direct.
    void* direct = nullptr;
    if (!has_rect && (fmt == FVSRN_OUTPUT_RGBA8 || fmt == FVSRN_OUTPUT_RGBA_F16)) {     // (the raw tensor's eight planes: 4.75 - 6.2 ms direct vs 4.91 staged)
        cudaPointerAttributes at{};
        if (cudaPointerGetAttributes(&at, out_host) == cudaSuccess && at.type == cudaMemoryTypeHost && at.devicePointer != nullptr) direct = at.devicePointer;
        else (void)cudaGetLastError();      // pageable memory: not an error, the staged path below handles it
#if defined(FVSRN_EXPERIMENTS) || defined(FVSRN_HOST_BANDS_ENV)
        if (std::getenv("FVSRN_HOST_STAGED")) direct = nullptr;
#endif
    }
    const int bands = (p->batch == 1 && !has_rect && !p->view_times_host && H >= 256 && net->aux_stream) ? 4 : 1;
    if (direct) {
        do_render(net, &q, reinterpret_cast<float*>(direct), s);
    } else if (bands == 1) {
        if (has_rect) CUDA_OK(cudaMemsetAsync(d_img, 0, img_b, s));
        do_render(net, &q, reinterpret_cast<float*>(d_img), s);
        CUDA_OK(cudaMemcpyAsync(out_host, d_img, img_b, cudaMemcpyDeviceToHost, s));
    } else {
        CUDA_OK(cudaEventRecord(net->ev_fork, s));
        CUDA_OK(cudaStreamWaitEvent(net->aux_stream, net->ev_fork, 0));
        // band heights in sixteenths of the frame, largest first: what is left exposed at the end is the LAST band's copy
        // (and its kernel's drain), so the bands shrink towards the end of the frame. A/B on one B200 (C3, fp32 colour, kernel
        // 4.67 ms; tools/ab_bands.sh): four equal bands 4.89 ms, 6/5/3/2 4.84, six or seven bands (5/4/3/2/1/1 ...) 4.98-5.01
        int pattern[16] = {6, 5, 3, 2}, npat = 4;
#if defined(FVSRN_EXPERIMENTS) || defined(FVSRN_HOST_BANDS_ENV)
        if (const char* e = std::getenv("FVSRN_HOST_BANDS")) {
            npat = 0;
            for (const char* c = e; *c && npat < 16; ++c) if (*c >= '1' && *c <= '9') pattern[npat++] = *c - '0';
        }
#endif
        int y_end[16], acc16 = 0, total16 = 0;
        for (int i = 0; i < npat; ++i) total16 += pattern[i];
        for (int i = 0; i < npat; ++i) { acc16 += pattern[i]; y_end[i] = i + 1 == npat ? int(H) : int((int64_t(H) * acc16 / total16 + 7) / 8 * 8); }
        int k = 0;
        for (int y0 = 0; k < npat && y0 < H; ++k) {
            const int h = std::min(y_end[k], int(H)) - y0;
            if (h <= 0) continue;
            cudaStream_t bs = (k & 1) ? net->aux_stream : s;
            q.tile_x0 = 0; q.tile_y0 = y0; q.tile_w = int(W); q.tile_h = h;
            do_render(net, &q, reinterpret_cast<float*>(d_img), bs);
            for (int c = 0; c < planes; ++c) {
                const size_t off = (size_t(c) * W * H + size_t(y0) * W) * elem;
                CUDA_OK(cudaMemcpyAsync(reinterpret_cast<char*>(out_host) + off, d_img + off, size_t(h) * W * elem, cudaMemcpyDeviceToHost, bs));
            }
            y0 += h;
        }
        CUDA_OK(cudaEventRecord(net->ev_join, net->aux_stream));
        CUDA_OK(cudaStreamWaitEvent(s, net->ev_join, 0));
    }
    CUDA_OK(cudaStreamSynchronize(s));
    return 0;
    FV_CATCH
}

int fvsrn_extract_color(const float* img_dev, float* out_dev, int batch, int height, int width, void* stream) {
    FV_TRY
    if (!img_dev || !out_dev) return fail("null device pointer");
    const long long hw = (long long)height * width, total = hw * 4 * batch;
    if (total <= 0) return 0;
    const int threads = 256;
    const int blocks = int(std::min<long long>((total + threads - 1) / threads, 148 * 16));
    extract_color_kernel<<<blocks, threads, 0, static_cast<cudaStream_t>(stream)>>>(img_dev, out_dev, hw, total);
    CUDA_OK(cudaGetLastError());
    return 0;
    FV_CATCH
}

int fvsrn_extract_color_ex(const float* img_dev, void* out_dev, int batch, int height, int width,
                           int channel_mode, int tonemap, float max_exposure, int output_format, void* stream) {
    FV_TRY
    if (!img_dev || !out_dev) return fail("null device pointer");
    if (output_format < FVSRN_OUTPUT_RGBA_F32 || output_format > FVSRN_OUTPUT_RGBA_F16) return fail("output format must be RGBA_F32, RGBA8 or RGBA_F16");
    if (channel_mode < 0 || channel_mode > FVSRN_CHANNEL_COLOR) return fail("unknown channel mode");
    if (tonemap && channel_mode == FVSRN_CHANNEL_COLOR && !(max_exposure > 0.f)) return fail("tonemapping needs a positive max_exposure");
    const long long hw = (long long)height * width, total = hw * batch;
    if (total <= 0) return 0;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    float* mm = nullptr;
    if (channel_mode == FVSRN_CHANNEL_DEPTH) {
        CUDA_OK(cudaMallocAsync(&mm, 2 * sizeof(float), s));
        depth_minmax_kernel<<<1, 1024, 0, s>>>(img_dev, hw, batch, mm);
        CUDA_OK(cudaGetLastError());
    }
    const int threads = 256;
    const int blocks = int(std::min<long long>((total + threads - 1) / threads, 148 * 16));
    extract_color_ex_kernel<<<blocks, threads, 0, s>>>(img_dev, out_dev, hw, batch, channel_mode, tonemap, max_exposure, output_format, mm);
    CUDA_OK(cudaGetLastError());
    if (mm) CUDA_OK(cudaFreeAsync(mm, s));
    return 0;
    FV_CATCH
}

// ---------------------------------------------------------------------------------------------------------------------
// fvsrn_extract_color_ex_noalloc: the same pass for callers that cannot allocate (CUDA graph capture, serving loops).
// fvsrn_extract_color_ex takes the two floats of its ChannelDepth min / max reduction from cudaMallocAsync; this variant keeps
// its scratch in a file-scope __device__ array instead: one {min, max, saw-NaN} triple per block of kNoAllocPixelsPerPartial
// pixels, then the reduced {min, max} in the first two floats of the tail. Every GPU that loads the library holds the
// array (kNoAllocScratchFloats * 4 = 192 KiB + 256 B) until exit. All calls share it: calls that overlap on the GPU (two
// streams, two host threads) race. Every word a call reads is written by that call first, so nothing has to be cleared.
} // extern "C"
namespace {
constexpr long long kNoAllocPixelsPerPartial = 4096;
constexpr long long kNoAllocMaxPartials = FVSRN_NOALLOC_MAX_PIXELS / kNoAllocPixelsPerPartial;      // 16384
constexpr long long kNoAllocScratchFloats = 3 * kNoAllocMaxPartials + 64;
static __device__ __align__(256) float g_noalloc_scratch[kNoAllocScratchFloats];

// device address of g_noalloc_scratch on the current device. In host code the array's name is only a stand-in for the
// symbol, not a device pointer: the address comes from cudaGetSymbolAddress, once per device, with that device current.
float* noalloc_scratch_ptr() {
    static std::mutex mu;
    static float* ptrs[64] = {};
    int dev = -1;
    CUDA_OK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) throw std::runtime_error("device index out of range");
    std::lock_guard<std::mutex> lock(mu);
    if (!ptrs[dev]) {
        float* p = nullptr;
        CUDA_OK(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_noalloc_scratch));
        if (!p) throw std::runtime_error("cudaGetSymbolAddress returned a null scratch address");
        ptrs[dev] = p;
    }
    return ptrs[dev];
}

// stage 1: block k reduces pixels [k * 4096, (k + 1) * 4096) of the depth plane -> partial[3k..3k+2] = {min, max, saw NaN}
__global__ void __launch_bounds__(256)
depth_partials_kernel(const float* __restrict__ img, long long hw, long long total, float* __restrict__ partial) {
    __shared__ float smin[8], smax[8];
    __shared__ int snan[8];
    float mn = FLT_MAX, mx = -FLT_MAX; int nan = 0;
    const long long first = (long long)blockIdx.x * kNoAllocPixelsPerPartial;
    for (long long i = first + threadIdx.x; i < first + kNoAllocPixelsPerPartial && i < total; i += blockDim.x) {
        const long long b = i / hw, px = i - b * hw;
        const float d = img[b * 8 * hw + 7 * hw + px];
        if (d != d) nan = 1; else { mn = fminf(mn, d); mx = fmaxf(mx, d); }
    }
    for (int o = 16; o > 0; o >>= 1) {
        mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, o)); mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        nan |= __shfl_xor_sync(0xffffffffu, nan, o);
    }
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    if (l == 0) { smin[w] = mn; smax[w] = mx; snan[w] = nan; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int k = 1; k < 8; ++k) { mn = fminf(mn, smin[k]); mx = fmaxf(mx, smax[k]); nan |= snan[k]; }
        partial[3 * blockIdx.x + 0] = mn; partial[3 * blockIdx.x + 1] = mx; partial[3 * blockIdx.x + 2] = nan ? 1.f : 0.f;
    }
}
// stage 2 (one block): the partials -> mm = {min, max}, NaN when any depth was NaN (torch.min / torch.max semantics)
__global__ void __launch_bounds__(1024)
depth_final_kernel(const float* __restrict__ partial, int n, float* __restrict__ mm) {
    __shared__ float smin[32], smax[32];
    __shared__ int snan[32];
    float mn = FLT_MAX, mx = -FLT_MAX; int nan = 0;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        mn = fminf(mn, partial[3 * i]); mx = fmaxf(mx, partial[3 * i + 1]); nan |= partial[3 * i + 2] != 0.f;
    }
    for (int o = 16; o > 0; o >>= 1) {
        mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, o)); mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        nan |= __shfl_xor_sync(0xffffffffu, nan, o);
    }
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    if (l == 0) { smin[w] = mn; smax[w] = mx; snan[w] = nan; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int k = 1; k < 32; ++k) { mn = fminf(mn, smin[k]); mx = fmaxf(mx, smax[k]); nan |= snan[k]; }
        mm[0] = nan ? __int_as_float(0x7fc00000) : mn; mm[1] = nan ? __int_as_float(0x7fc00000) : mx;
    }
}
} // namespace
extern "C" {

int fvsrn_extract_color_ex_noalloc(const float* img_dev, void* out_dev, int batch, int height, int width,
                                   int channel_mode, int tonemap, float max_exposure, int output_format, void* stream) {
    FV_TRY
    if (!img_dev || !out_dev) return fail("null device pointer");
    if (output_format < FVSRN_OUTPUT_RGBA_F32 || output_format > FVSRN_OUTPUT_RGBA_F16) return fail("output format must be RGBA_F32, RGBA8 or RGBA_F16");
    if (channel_mode < 0 || channel_mode > FVSRN_CHANNEL_COLOR) return fail("unknown channel mode");
    if (tonemap && channel_mode == FVSRN_CHANNEL_COLOR && !(max_exposure > 0.f)) return fail("tonemapping needs a positive max_exposure");
    if (batch < 0 || height < 0 || width < 0) return fail("negative image size");
    const long long hw = (long long)height * width, total = hw * batch;
    if (total > FVSRN_NOALLOC_MAX_PIXELS)
        return fail("fvsrn_extract_color_ex_noalloc: batch * height * width = " + std::to_string(total) + " pixels exceeds the static scratch limit of " +
                    std::to_string((long long)FVSRN_NOALLOC_MAX_PIXELS) + " (FVSRN_NOALLOC_MAX_PIXELS); use fvsrn_extract_color_ex");
    if (total <= 0) return 0;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    float* scratch = noalloc_scratch_ptr();
    float* mm = scratch + 3 * kNoAllocMaxPartials;      // the 64-float tail holds the reduced {min, max}
    if (channel_mode == FVSRN_CHANNEL_DEPTH) {
        const int parts = int((total + kNoAllocPixelsPerPartial - 1) / kNoAllocPixelsPerPartial);
        depth_partials_kernel<<<parts, 256, 0, s>>>(img_dev, hw, total, scratch);
        CUDA_OK(cudaGetLastError());
        depth_final_kernel<<<1, 1024, 0, s>>>(scratch, parts, mm);
        CUDA_OK(cudaGetLastError());
    }
    const int threads = 256;
    const int blocks = int(std::min<long long>((total + threads - 1) / threads, 148 * 16));
    extract_color_ex_kernel<<<blocks, threads, 0, s>>>(img_dev, out_dev, hw, batch, channel_mode, tonemap, max_exposure, output_format, mm);
    CUDA_OK(cudaGetLastError());
    return 0;
    FV_CATCH
}

// ---------------------------------------------------------------------------------------------------------------------
// Peer frame buffers (multi-GPU, one process per GPU): a device buffer on the collecting rank that the other ranks' DVR kernels
// write their pixels into directly over NVLink (CUDA IPC mapping) - the render and the "gather" become one kernel per rank.
int fvsrn_peer_buffer_create(size_t bytes, void** dev_ptr, unsigned char handle[64]) {
    FV_TRY
    if (!dev_ptr || !handle || bytes == 0) return fail("null argument");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "CUDA IPC handle size");
    void* p = nullptr;
    CUDA_OK(cudaMalloc(&p, bytes));             // its own allocation: an IPC handle names a whole cudaMalloc block
    cudaIpcMemHandle_t h;
    const cudaError_t e = cudaIpcGetMemHandle(&h, p);
    if (e != cudaSuccess) { cudaFree(p); return fail(std::string("cudaIpcGetMemHandle failed: ") + cudaGetErrorString(e)); }
    CUDA_OK(cudaMemset(p, 0, bytes));
    std::memcpy(handle, &h, 64);
    *dev_ptr = p;
    return 0;
    FV_CATCH
}
int fvsrn_peer_buffer_open(const unsigned char handle[64], void** dev_ptr) {
    FV_TRY
    if (!dev_ptr || !handle) return fail("null argument");
    cudaIpcMemHandle_t h;
    std::memcpy(&h, handle, 64);
    void* p = nullptr;
    const cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) return fail(std::string("cudaIpcOpenMemHandle failed (peer access between the two GPUs is required): ") + cudaGetErrorString(e));
    *dev_ptr = p;
    return 0;
    FV_CATCH
}
int fvsrn_peer_buffer_close(void* dev_ptr) {
    FV_TRY
    if (dev_ptr) CUDA_OK(cudaIpcCloseMemHandle(dev_ptr));
    return 0;
    FV_CATCH
}
int fvsrn_peer_buffer_free(void* dev_ptr) {
    FV_TRY
    if (dev_ptr) CUDA_OK(cudaFree(dev_ptr));
    return 0;
    FV_CATCH
}

int fvsrn_generate_rays(const float* camera_matrix_dev, int batch, float fov_y_radians, int width, int height,
                        float* ray_start_dev, float* ray_dir_dev, void* stream) {
    FV_TRY
    if (!camera_matrix_dev || !ray_start_dev || !ray_dir_dev) return fail("null device pointer");
    if (batch <= 0 || width <= 0 || height <= 0) return fail("image size and batch must be positive");
    CUDA_OK(launch_generate_rays(camera_matrix_dev, fov_y_radians, float(double(width) / double(height)), batch, height, width,
                                 ray_start_dev, ray_dir_dev, static_cast<cudaStream_t>(stream)));
    return 0;
    FV_CATCH
}

int fvsrn_evaluate_tf(int tf_kind, const float* tf_dev, int tf_resolution, const float* densities_dev, float* colors_dev,
                      int64_t n, float density_min, float density_max, float stepsize, void* stream) {
    FV_TRY
    if (n < 0) return fail("negative sample count");
    if (n == 0) return 0;
    if (!tf_dev || !densities_dev || !colors_dev) return fail("null device pointer");
    if (tf_kind < 0 || tf_kind > 2) return fail("unknown transfer function kind");
    if (tf_kind != FVSRN_TF_IDENTITY && tf_resolution < 2) return fail("transfer function needs at least 2 entries");
    CUDA_OK(launch_evaluate_tf(tf_kind, tf_dev, tf_resolution, densities_dev, colors_dev, (long long)n, density_min, density_max, stepsize,
                               static_cast<cudaStream_t>(stream)));
    return 0;
    FV_CATCH
}

int fvsrn_query(fvsrn_network* net, int kernel, fvsrn_query_info* info) {
    FV_TRY
    if (!net) return fail("null network handle");
    std::lock_guard<std::mutex> lock(net->mu);
    DeviceGuard guard(net->device);
    require_prepared(net);
    if (!info) return fail("null argument");
    const bool tc = net->kernel_mode == FVSRN_KERNEL_TCGEN05 && net->net.output_mode <= OUT_RGBO_DIRECT &&
                    (kernel != 0 || net->gradient_mode == FVSRN_GRADIENT_OFF_OR_DIRECT);
    const NetParams P = make_net_params(net, net->net.current_time, net->net.current_ensemble, false, tc);
    LaunchCfg cfg; cfg.sm_count = net->sm_count; cfg.query_only = 1;
    launch_fn fn = pick_launcher(net->H16);
    if (!fn) return fail("no kernel for this hidden width");
    DvrParams dq{}; dq.tf_kind = 0; dq.tf_R = 8;   // a typical piecewise TF: the DVR kernels size their shared memory from it
    dq.tw = 1024; dq.th = 1024; dq.nb = 1;
    if (net->gradient_mode == FVSRN_GRADIENT_ADJOINT_METHOD && net->kernel_mode == FVSRN_KERNEL_REFERENCE_EXACT && kernel == 0) CUDA_OK(launch_exact_adjoint(net->H16, net->GC16, P, dq, &cfg, nullptr));
    else if (net->kernel_mode == FVSRN_KERNEL_REFERENCE_EXACT && kernel == 0) CUDA_OK(launch_exact(net->H16, net->GC16, net->dir, P, dq, &cfg, nullptr));
    else CUDA_OK(fn((kernel == 0 ? KIND_DVR : KIND_EVAL) + (tc ? 2 : 0), net->GC16, net->dir, P, kernel == 0 ? &dq : nullptr, &cfg, nullptr));
    info->device = net->device; info->sm_count = net->sm_count;
    info->block_threads = cfg.threads; info->blocks_per_sm = cfg.blocks_per_sm; info->grid_blocks = cfg.grid;
    info->smem_bytes_per_block = cfg.smem; info->regs_per_thread = cfg.regs;
    info->rays_per_warp_tile_x = cfg.tile_w ? cfg.tile_w : (tc ? kTcTileW : kTileW);
    info->rays_per_warp_tile_y = cfg.tile_h ? cfg.tile_h : (tc ? kTcTileH : kTileH);
    info->kernel_launches = net->launches;
    return 0;
    FV_CATCH
}

} // extern "C"
