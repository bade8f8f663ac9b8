// Host-side model of a scene representation network (SRN) and its .volnet serialisation.
// B200-native counterpart of renderer::SceneNetwork (reference renderer/volume_interpolation_network.{h,cpp}).
// Pure host C++ (no CUDA calls) so that load/save/validate work on a machine without a GPU.
#pragma once
#include <cstdint>
#include <cstring>
#include <string>
#include <vector>
#include <stdexcept>

namespace fvsrn {

// fp16 <-> fp32 on the host, round-to-nearest-even (same as __float2half on the reference host path)
uint16_t f32_to_f16(float f);
float f16_to_f32(uint16_t h);

enum Activation { ACT_RELU = 0, ACT_SINE = 1, ACT_SNAKE = 2, ACT_SNAKEALT = 3, ACT_SIGMOID = 4, ACT_NONE = 5, NUM_ACTIVATIONS = 6 };
enum OutputMode {
    OUT_DENSITY = 0, OUT_DENSITY_DIRECT = 1, OUT_RGBO = 2, OUT_RGBO_DIRECT = 3,
    OUT_DENSITYGRAD = 4, OUT_DENSITYGRAD_DIRECT = 5, OUT_DENSITYGRAD_CUBIC = 6,
    OUT_DENSITYCURV = 7, OUT_DENSITYCURV_DIRECT = 8, NUM_OUTPUT_MODES = 9
};
enum GridEncoding { GRID_FLOAT = 0, GRID_BYTE_LINEAR = 1, GRID_BYTE_GAUSSIAN = 2 };

extern const char* const kActivationNames[NUM_ACTIVATIONS];
extern const char* const kOutputModeNames[NUM_OUTPUT_MODES];
extern const int kOutputModeChannelsIn[NUM_OUTPUT_MODES];
extern const int kOutputModeChannelsOut[NUM_OUTPUT_MODES];
int activation_from_string(const std::string& s);   // throws
int output_mode_from_string(const std::string& s);  // throws

// One linear layer exactly as the .volnet stores it (i.e. AFTER SceneNetwork::addLayer's re-layout,
// reference volume_interpolation_network.cpp:806-894).
struct Layer {
    int cin = 0, cout = 0;
    std::vector<uint16_t> weights; // fp16 bits
    std::vector<uint16_t> bias;    // fp16 bits
    int activation = ACT_NONE;
    float activation_param = 1.f;
    bool transposed = false;       // stored as [cout + channelsOut*cin] (first w/o Fourier, or last layer)
    // canonical accessor W[co][ci]
    uint16_t w(int co, int ci) const { return transposed ? weights[co + cout * ci] : weights[co * cin + ci]; }
};

// One latent grid as stored (reference LatentGrid, :290-334, 564-614)
struct LatentGrid {
    int encoding = GRID_FLOAT;
    int C = 0, Z = 0, Y = 0, X = 0;
    std::vector<char> data;           // (C/4, Z, Y, X, 4) of float32 or uint8
    std::vector<float> offset_or_mean, scale_or_std;
    size_t bytes_per_entry() const { return encoding == GRID_FLOAT ? 4 : 1; }
    size_t idx(int cHigh, int z, int y, int x, int cLow) const {
        return size_t(cLow) + 4 * (size_t(x) + X * (size_t(y) + Y * (size_t(z) + size_t(Z) * cHigh)));
    }
    bool valid() const { return C > 0 && C % 4 == 0 && Z > 0 && Y > 0 && X > 0 && data.size() == bytes_per_entry() * size_t(C) * Z * Y * X; }
    // decoded value of channel c at voxel (z,y,x) in fp32 (reference renderer_volume_tensorcores.cuh:325-384)
    float decoded(int c, int z, int y, int x) const;
    // build from a (C,Z,Y,X) fp32 tensor; returns mean abs encoding error
    double from_dense(const float* grid, int C_, int Z_, int Y_, int X_, int encoding_);
};

struct Network {
    // InputParametrization
    bool has_time = false, has_direction = false, direction_in_fourier = false;
    int num_fourier = 0;
    std::vector<uint16_t> fourier; // index = cout + num_fourier*cin, already x 2pi
    // OutputParametrization
    int output_mode = OUT_DENSITY;
    std::vector<Layer> layers;
    float box_min[3] = {-5.f, -5.f, -5.f};  // reference default, :797-804
    float box_size[3] = {1.f, 1.f, 1.f};
    // LatentGridTimeAndEnsemble
    bool has_latent_grid = false;
    int time_min = 0, time_num = 0, time_step = 1, ensemble_min = 0, ensemble_num = 0;
    std::vector<LatentGrid> time_grids, ensemble_grids;
    float current_time = 0.f;
    int current_ensemble = 0;

    int input_channels_out() const; // InputParametrization::channelsOut :31-47
    int latent_channels() const;
    int time_channels() const { return time_grids.empty() ? 0 : time_grids[0].C; }
    int ensemble_channels() const { return ensemble_grids.empty() ? 0 : ensemble_grids[0].C; }
    int time_max_inclusive() const { return time_min + (time_num - 1) * time_step; }
    int ensemble_max_inclusive() const { return ensemble_min + ensemble_num - 1; }
    float interpolate_time(float t) const;  // absolute time -> keyframe index (float), :360-364
    int interpolate_ensemble(int e) const;

    void add_layer(const Layer& canonical_row_major); // applies the reference's re-layout
    bool valid(std::string* why = nullptr) const;
    int num_parameters() const;
    int compute_max_warps(bool only_shared, bool adjoint) const;
    void set_time_and_ensemble(float time, int ensemble);

    static Network load(const uint8_t* bytes, size_t n);
    std::vector<uint8_t> save() const;
};

} // namespace fvsrn
