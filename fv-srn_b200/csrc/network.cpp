// .volnet I/O, layer re-layout and validity rules of the SRN host model.
// Behaviour follows reference renderer/volume_interpolation_network.cpp (file:line cited per function);
// the code is written from the format description, not translated.
#include "network.hpp"
#include <algorithm>
#include <cmath>
#include <cstdio>

namespace fvsrn {

// ---------------------------------------------------------------- fp16 helpers
uint16_t f32_to_f16(float f) {
    uint32_t x; std::memcpy(&x, &f, 4);
    const uint32_t sign = (x >> 16) & 0x8000u;
    const uint32_t absx = x & 0x7fffffffu;
    if (absx >= 0x7f800000u) { // inf / nan
        return uint16_t(sign | 0x7c00u | ((absx > 0x7f800000u) ? 0x200u : 0u));
    }
    if (absx >= 0x477ff000u) { // rounds to >= 65520 -> inf
        return uint16_t(sign | 0x7c00u);
    }
    if (absx < 0x33000001u) { // < 2^-25 (or exactly 2^-25 -> ties to even = 0)
        return uint16_t(sign);
    }
    int exp = int(absx >> 23) - 127;
    uint32_t mant = (absx & 0x7fffffu) | 0x800000u; // 24 bits
    int shift; int hexp;
    if (exp < -14) { shift = 13 + (-14 - exp); hexp = 0; } // subnormal half
    else { shift = 13; hexp = exp + 15; }
    uint32_t hm = mant >> shift;
    const uint32_t rem = mant & ((1u << shift) - 1u);
    const uint32_t half = 1u << (shift - 1);
    if (rem > half || (rem == half && (hm & 1u))) hm++;
    uint32_t out;
    if (hexp == 0) out = hm;                          // subnormal (may carry into exponent 1: correct)
    else out = (uint32_t(hexp) << 10) + (hm - 0x400u); // carry propagates into exponent: correct
    return uint16_t(sign | out);
}

float f16_to_f32(uint16_t h) {
    const uint32_t sign = uint32_t(h & 0x8000u) << 16;
    const uint32_t exp = (h >> 10) & 0x1fu;
    const uint32_t mant = h & 0x3ffu;
    uint32_t out;
    if (exp == 0) {
        if (mant == 0) out = sign;
        else { // subnormal
            int e = -1; uint32_t m = mant;
            do { e++; m <<= 1; } while ((m & 0x400u) == 0);
            out = sign | uint32_t(127 - 15 - e) << 23 | ((m & 0x3ffu) << 13);
        }
    } else if (exp == 31) out = sign | 0x7f800000u | (mant << 13);
    else out = sign | ((exp + 112u) << 23) | (mant << 13);
    float f; std::memcpy(&f, &out, 4); return f;
}

// ---------------------------------------------------------------- names
const char* const kActivationNames[NUM_ACTIVATIONS] = {"ReLU", "Sine", "Snake", "SnakeAlt", "Sigmoid", "None"};
const char* const kOutputModeNames[NUM_OUTPUT_MODES] = {
    "density", "density:direct", "rgbo", "rgbo:direct", "densitygrad", "densitygrad:direct",
    "densitygrad:cubic", "densitycurvature", "densitycurvature:direct"};
const int kOutputModeChannelsIn[NUM_OUTPUT_MODES] = {1, 1, 4, 4, 4, 4, 4, 6, 6};
const int kOutputModeChannelsOut[NUM_OUTPUT_MODES] = {1, 1, 4, 4, 1, 1, 1, 1, 1};

int activation_from_string(const std::string& s) {
    for (int i = 0; i < NUM_ACTIVATIONS; ++i) if (s == kActivationNames[i]) return i;
    throw std::runtime_error("No activation found matching string " + s);
}
int output_mode_from_string(const std::string& s) {
    for (int i = 0; i < NUM_OUTPUT_MODES; ++i) if (s == kOutputModeNames[i]) return i;
    throw std::runtime_error("No output mode found matching string " + s);
}

// ---------------------------------------------------------------- byte stream helpers
namespace {
struct Reader {
    const uint8_t* p; size_t n, pos = 0;
    void raw(void* dst, size_t k) {
        if (pos + k > n) throw std::runtime_error("volnet: unexpected end of data");
        std::memcpy(dst, p + pos, k); pos += k;
    }
    template <typename T> T get() { T v; raw(&v, sizeof(T)); return v; }
    std::string str() {
        int l = get<int32_t>();
        if (l < 0 || size_t(l) > n - pos) throw std::runtime_error("volnet: bad string length");
        std::string s(size_t(l), '\0'); raw(&s[0], size_t(l)); return s;
    }
};
struct Writer {
    std::vector<uint8_t> buf;
    void raw(const void* src, size_t k) { const uint8_t* s = (const uint8_t*)src; buf.insert(buf.end(), s, s + k); }
    template <typename T> void put(T v) { raw(&v, sizeof(T)); }
    void str(const std::string& s) { put<int32_t>(int32_t(s.size())); raw(s.data(), s.size()); }
};

// inverse error function (double), used only for the byte-gaussian grid decode
double erfinv_d(double x) {
    if (x <= -1) return -INFINITY;
    if (x >= 1) return INFINITY;
    // initial guess (Winitzki) + two Newton steps on erf
    const double a = 0.147;
    const double ln = std::log(1 - x * x);
    const double t = 2 / (M_PI * a) + ln / 2;
    double y = std::copysign(std::sqrt(std::sqrt(t * t - ln / a) - t), x);
    for (int i = 0; i < 3; ++i) {
        const double err = std::erf(y) - x;
        y -= err / (2 / std::sqrt(M_PI) * std::exp(-y * y));
    }
    return y;
}
} // namespace

// ---------------------------------------------------------------- LatentGrid
float LatentGrid::decoded(int c, int z, int y, int x) const {
    const size_t i = idx(c / 4, z, y, x, c % 4);
    if (encoding == GRID_FLOAT) { float v; std::memcpy(&v, data.data() + 4 * i, 4); return v; }
    const float u = float(uint8_t(data[i])) / 255.0f; // what an 8-bit unorm texture returns
    if (encoding == GRID_BYTE_LINEAR) return offset_or_mean[c] + u * scale_or_std[c];
    // byte-gaussian: mean + std * sqrt2 * erfinv((2-eps)(u-0.5)), reference renderer_volume_tensorcores.cuh:365-383
    const float t = float(1.4142135623730951 * erfinv_d(double((2.0f - 1e-4f) * (u - 0.5f))));
    return offset_or_mean[c] + t * scale_or_std[c];
}

double LatentGrid::from_dense(const float* g, int C_, int Z_, int Y_, int X_, int encoding_) {
    C = C_; Z = Z_; Y = Y_; X = X_; encoding = encoding_;
    const size_t vox = size_t(Z) * Y * X;
    if (C <= 0 || C % 4 != 0) throw std::runtime_error("LatentGrid: channel count must be a positive multiple of 4");
    offset_or_mean.clear(); scale_or_std.clear();
    double err = 0;
    if (encoding == GRID_FLOAT) { // reference initEncodingFloat :319-334
        data.assign(4 * size_t(C) * vox, 0);
        float* d = reinterpret_cast<float*>(data.data());
        for (int c = 0; c < C; ++c) for (int z = 0; z < Z; ++z) for (int y = 0; y < Y; ++y) for (int x = 0; x < X; ++x)
            d[idx(c / 4, z, y, x, c % 4)] = g[(size_t(c) * Z + z) * Y * X + size_t(y) * X + x];
        return 0;
    }
    data.assign(size_t(C) * vox, 0);
    offset_or_mean.resize(C); scale_or_std.resize(C);
    for (int c = 0; c < C; ++c) {
        const float* gc = g + size_t(c) * vox;
        if (encoding == GRID_BYTE_LINEAR) { // reference initEncodingByteLinear :336-381
            float mn = gc[0], mx = gc[0];
            for (size_t i = 1; i < vox; ++i) { mn = std::min(mn, gc[i]); mx = std::max(mx, gc[i]); }
            offset_or_mean[c] = mn; scale_or_std[c] = mx - mn;
            const float inv = 1.0f / std::max(1e-5f, mx - mn);
            for (int z = 0; z < Z; ++z) for (int y = 0; y < Y; ++y) for (int x = 0; x < X; ++x) {
                const float v = gc[(size_t(z) * Y + y) * X + x];
                const int q = std::max(0, std::min(255, int(std::roundf(255 * ((v - mn) * inv)))));
                data[idx(c / 4, z, y, x, c % 4)] = char(uint8_t(q));
                err += std::fabs(v - (mn + q / 255.0f * (mx - mn)));
            }
        } else if (encoding == GRID_BYTE_GAUSSIAN) { // reference initEncodingByteGaussian :383-432
            double s = 0, s2 = 0;
            for (size_t i = 0; i < vox; ++i) s += gc[i];
            const double mean = s / double(vox);
            for (size_t i = 0; i < vox; ++i) s2 += (gc[i] - mean) * (gc[i] - mean);
            const float sd = float(std::sqrt(s2 / double(std::max<size_t>(1, vox - 1)))); // torch::std is unbiased
            offset_or_mean[c] = float(mean); scale_or_std[c] = sd;
            const float inv = 1.0f / std::max(1e-5f, sd);
            for (int z = 0; z < Z; ++z) for (int y = 0; y < Y; ++y) for (int x = 0; x < X; ++x) {
                const float v = gc[(size_t(z) * Y + y) * X + x];
                const float th = 0.5f * (1 + std::erf((v - float(mean)) * inv * 0.70710678f));
                const int q = std::max(0, std::min(255, int(std::roundf(255 * th))));
                data[idx(c / 4, z, y, x, c % 4)] = char(uint8_t(q));
                err += std::fabs(v - decoded(c, z, y, x));
            }
        } else throw std::runtime_error("Unsupported encoding");
    }
    return err / double(size_t(C) * vox);
}

// ---------------------------------------------------------------- Network
int Network::input_channels_out() const { // reference :31-41
    if (num_fourier > 0) return 4 + (has_direction ? 4 : 0) + 2 * num_fourier;
    return 3 + (has_direction ? 3 : 0);
}
int Network::latent_channels() const { return has_latent_grid ? time_channels() + ensemble_channels() : 0; }
float Network::interpolate_time(float t) const { // reference volume_interpolation_network.h:360-364
    const float v = (t - float(time_min)) / float(time_step);
    return std::min(std::max(v, 0.f), float(time_num - 1));
}
int Network::interpolate_ensemble(int e) const { return std::min(std::max(e - ensemble_min, 0), ensemble_num - 1); }

void Network::set_time_and_ensemble(float time, int ensemble) { // reference :923-938 (silently clamps)
    if (!has_latent_grid) return;
    current_time = std::min(std::max(time, float(time_min)), float(time_max_inclusive()));
    current_ensemble = std::min(std::max(ensemble, ensemble_min), ensemble_max_inclusive());
}

// SceneNetwork::addLayer, reference :806-894. Input: canonical row-major W[cout][cin].
void Network::add_layer(const Layer& in) {
    Layer l = in; l.transposed = false;
    const int co = in.cout, ci = in.cin;
    auto W = [&](int o, int i) { return in.weights[size_t(o) * ci + i]; };
    if (layers.empty() && num_fourier > 0) {
        // first layer: insert the skipped (zero) input columns so that inputs are
        // [px,py,pz,time|0,(dx,dy,dz,0),cos..,sin..,latent..]
        int nci = ci;
        std::vector<uint16_t> w;
        if (!has_time) {
            if (has_direction) {
                nci = ci + 2; w.assign(size_t(nci) * co, 0);
                for (int o = 0; o < co; ++o) {
                    for (int i = 0; i < 3; ++i) w[size_t(o) * nci + i] = W(o, i);
                    for (int i = 0; i < 3; ++i) w[size_t(o) * nci + i + 4] = W(o, i + 3);
                    for (int i = 0; i < ci - 6; ++i) w[size_t(o) * nci + i + 8] = W(o, i + 6);
                }
            } else {
                nci = ci + 1; w.assign(size_t(nci) * co, 0);
                for (int o = 0; o < co; ++o) {
                    for (int i = 0; i < 3; ++i) w[size_t(o) * nci + i] = W(o, i);
                    for (int i = 0; i < ci - 3; ++i) w[size_t(o) * nci + i + 4] = W(o, i + 3);
                }
            }
        } else if (has_direction) {
            nci = ci + 1; w.assign(size_t(nci) * co, 0);
            for (int o = 0; o < co; ++o) {
                for (int i = 0; i < 4; ++i) w[size_t(o) * nci + i] = W(o, i);
                for (int i = 0; i < 3; ++i) w[size_t(o) * nci + i + 4] = W(o, i + 4);
                // NOTE: the reference copies (cin-6) trailing inputs starting at old index 7 (:866-869), i.e. it
                // reads one element past the row for the last one; we copy the (cin-7) that exist.
                for (int i = 0; i < ci - 7; ++i) w[size_t(o) * nci + i + 8] = W(o, i + 7);
            }
        } else {
            w = in.weights;
        }
        l.cin = nci; l.weights = w;
    } else if (ci < 16 || co < 16) {
        std::vector<uint16_t> w(in.weights.size());
        for (int o = 0; o < co; ++o) for (int i = 0; i < ci; ++i) w[o + size_t(co) * i] = W(o, i);
        l.weights = w; l.transposed = true;
    }
    layers.push_back(l);
}

bool Network::valid(std::string* why) const { // reference :940-985 (+ InputParametrization::valid :43-67)
    auto fail = [&](const std::string& s) { if (why) *why = s; return false; };
    if (direction_in_fourier && !has_direction) return fail("useDirectionInFourierFeatures requires hasDirection");
    const int fc = direction_in_fourier ? 6 : 3;
    if (int(fourier.size()) != num_fourier * fc) return fail("fourier matrix size does not match numFourierFeatures");
    if (num_fourier % 2 != 0) return fail("The number of fourier features must be divisible by 2");
    if (has_latent_grid) {
        for (const auto& g : time_grids) if (!g.valid() || g.C % 16 != 0) return fail("LatentGrid is invalid");
        for (const auto& g : ensemble_grids) if (!g.valid() || g.C % 16 != 0) return fail("LatentGrid is invalid");
        if (int(time_grids.size()) != time_num || int(ensemble_grids.size()) != ensemble_num) return fail("LatentGrid count mismatch");
        if (time_grids.empty() && ensemble_grids.empty()) return fail("LatentGrid has no grids");
        for (const auto& g : time_grids) if (g.C != time_grids[0].C || g.encoding != time_grids[0].encoding) return fail("time grids differ in channels/encoding");
        for (const auto& g : ensemble_grids) if (g.C != ensemble_grids[0].C) return fail("ensemble grids differ in channels");
        if (num_fourier == 0) return fail("Currently, LatentGrid requires fourier features as well");
    }
    if (layers.empty()) return fail("no layers");
    int cur = input_channels_out() + latent_channels();
    for (size_t i = 0; i < layers.size(); ++i) {
        const Layer& l = layers[i];
        if (l.cin != cur) {
            char b[160]; std::snprintf(b, sizeof b, "Invalid input channels at hidden layer %zu, expected %d, got %d", i, cur, l.cin);
            return fail(b);
        }
        const bool last = (i + 1 == layers.size());
        if (l.weights.size() != size_t(l.cin) * l.cout || int(l.bias.size()) != l.cout || (!last && l.cout % 4 != 0))
            return fail("Invalid hidden layer, weights and bias don't match or are not a multiple of 4");
        cur = l.cout;
    }
    if (cur != kOutputModeChannelsIn[output_mode]) return fail("Output channels from the hidden layers don't match the output parametrization");
    return true;
}

int Network::num_parameters() const { // reference :1043-1055
    int n = num_fourier ? int(fourier.size()) : 0;
    for (const auto& l : layers) n += int(l.weights.size() + l.bias.size());
    return n;
}

int Network::compute_max_warps(bool only_shared, bool adjoint) const { // reference :987-1041 (48 KiB static smem model)
    const int maxShared = 48 * 1024, maxConstant = 16 * 1024;
    int numShared = 0, numConst = 0;
    if (num_fourier) numConst += int(fourier.size());
    int maxChannels = input_channels_out();
    for (const auto& l : layers) {
        if (l.cin < 16 || l.cout < 16) numConst += int(l.weights.size() + l.bias.size());
        else numShared += int(l.weights.size() + l.bias.size());
        maxChannels = std::max(maxChannels, l.cout);
    }
    if (only_shared) { numShared += numConst; numConst = 0; }
    int perThread = maxChannels;
    if (adjoint) perThread += (int(layers.size()) - 1) * maxChannels;
    numShared *= 2; numConst *= 2; perThread *= 2;
    if (numConst > maxConstant) return -1;
    const int w = int(std::floor((maxShared - numShared) / float(perThread * 32)));
    return w <= 0 ? -1 : w;
}

// ---------------------------------------------------------------- .volnet
static LatentGrid load_grid(Reader& r) { // reference LatentGrid::load :564-593
    const int version = r.get<int32_t>();
    if (version != 1) throw std::runtime_error("Unknown version for LatentGrid " + std::to_string(version));
    LatentGrid g;
    g.encoding = r.get<int32_t>(); g.C = r.get<int32_t>(); g.Z = r.get<int32_t>(); g.Y = r.get<int32_t>(); g.X = r.get<int32_t>();
    if (g.encoding < 0 || g.encoding > 2 || g.C <= 0 || g.Z <= 0 || g.Y <= 0 || g.X <= 0) throw std::runtime_error("volnet: bad LatentGrid header");
    const size_t bytes = g.bytes_per_entry() * size_t(g.C) * g.Z * g.Y * g.X;
    if (bytes > r.n - r.pos) throw std::runtime_error("volnet: unexpected end of data");
    g.data.resize(bytes); r.raw(g.data.data(), bytes);
    if (g.encoding != GRID_FLOAT) {
        g.offset_or_mean.resize(g.C); g.scale_or_std.resize(g.C);
        r.raw(g.offset_or_mean.data(), 4 * size_t(g.C)); r.raw(g.scale_or_std.data(), 4 * size_t(g.C));
    }
    return g;
}
static void save_grid(Writer& w, const LatentGrid& g) { // reference :595-614
    w.put<int32_t>(1); w.put<int32_t>(g.encoding); w.put<int32_t>(g.C); w.put<int32_t>(g.Z); w.put<int32_t>(g.Y); w.put<int32_t>(g.X);
    w.raw(g.data.data(), g.data.size());
    if (g.encoding != GRID_FLOAT) { w.raw(g.offset_or_mean.data(), 4 * size_t(g.C)); w.raw(g.scale_or_std.data(), 4 * size_t(g.C)); }
}

Network Network::load(const uint8_t* bytes, size_t n) { // reference SceneNetwork::load :1059-1086
    Reader r{bytes, n};
    Network net;
    const int version = r.get<int32_t>();
    if (version > 2 || version < 1) throw std::runtime_error("Unknown version for SceneNetwork " + std::to_string(version));
    { // InputParametrization::load :70-114
        const int v = r.get<int32_t>();
        if (v < 1 || v > 3) throw std::runtime_error("Unknown version for InputParametrization " + std::to_string(v));
        if (v == 3) net.has_time = r.get<uint8_t>() != 0;
        net.has_direction = r.get<uint8_t>() != 0;
        net.num_fourier = r.get<int32_t>();
        net.direction_in_fourier = (v >= 2) ? (r.get<uint8_t>() != 0) : false;
        const int fc = net.direction_in_fourier ? 6 : 3;
        if (net.num_fourier < 0 || size_t(net.num_fourier) * fc * 2 > n) throw std::runtime_error("volnet: bad fourier count");
        net.fourier.resize(size_t(net.num_fourier) * fc);
        if (net.num_fourier > 0) r.raw(net.fourier.data(), 2 * net.fourier.size());
    }
    { // OutputParametrization::load :205-215
        const int v = r.get<int32_t>();
        if (v != 1) throw std::runtime_error("Unknown version for OutputParametrization " + std::to_string(v));
        net.output_mode = output_mode_from_string(r.str());
    }
    const int numLayers = r.get<int32_t>();
    if (numLayers < 0 || numLayers > 4096) throw std::runtime_error("volnet: bad layer count");
    for (int i = 0; i < numLayers; ++i) { // Layer::load :250-272
        const int v = r.get<int32_t>();
        if (v != 1 && v != 2) throw std::runtime_error("Unknown version for Layer " + std::to_string(v));
        Layer l;
        l.cout = r.get<int32_t>(); l.cin = r.get<int32_t>();
        if (l.cout <= 0 || l.cin <= 0 || size_t(l.cout) * l.cin * 2 > n) throw std::runtime_error("volnet: bad layer size");
        l.weights.resize(size_t(l.cout) * l.cin); l.bias.resize(l.cout);
        r.raw(l.weights.data(), 2 * l.weights.size()); r.raw(l.bias.data(), 2 * l.bias.size());
        l.activation = activation_from_string(r.str());
        l.activation_param = (v == 2) ? r.get<float>() : 1.f;
        // stored layout is implied by addLayer's rule (:806-894)
        l.transposed = !(i == 0 && net.num_fourier > 0) && (l.cin < 16 || l.cout < 16);
        net.layers.push_back(l);
    }
    r.raw(net.box_min, 12); r.raw(net.box_size, 12);
    if (version == 2) {
        const uint8_t hasGrid = r.get<uint8_t>();
        if (hasGrid > 0) { // LatentGridTimeAndEnsemble::load :758-780
            const int v = r.get<int32_t>();
            if (v > 1) throw std::runtime_error("Unknown version for LatentGridTimeAndEnsemble " + std::to_string(v));
            net.has_latent_grid = true;
            net.time_min = r.get<int32_t>(); net.time_num = r.get<int32_t>(); net.time_step = r.get<int32_t>();
            net.ensemble_min = r.get<int32_t>(); net.ensemble_num = r.get<int32_t>();
            if (net.time_num < 0 || net.ensemble_num < 0 || net.time_num > 65536 || net.ensemble_num > 65536) throw std::runtime_error("volnet: bad grid counts");
            for (int i = 0; i < net.time_num; ++i) net.time_grids.push_back(load_grid(r));
            for (int i = 0; i < net.ensemble_num; ++i) net.ensemble_grids.push_back(load_grid(r));
        }
    }
    return net;
}

std::vector<uint8_t> Network::save() const { // reference SceneNetwork::save :1088-1104 (always writes the newest versions)
    std::string why;
    if (!valid(&why)) throw std::runtime_error("scene network is not valid, cannot save: " + why);
    Writer w;
    w.put<int32_t>(2);
    w.put<int32_t>(3); w.put<uint8_t>(has_time); w.put<uint8_t>(has_direction); w.put<int32_t>(num_fourier); w.put<uint8_t>(direction_in_fourier);
    if (num_fourier > 0) w.raw(fourier.data(), 2 * fourier.size());
    w.put<int32_t>(1); w.str(kOutputModeNames[output_mode]);
    w.put<int32_t>(int32_t(layers.size()));
    for (const auto& l : layers) {
        w.put<int32_t>(2); w.put<int32_t>(l.cout); w.put<int32_t>(l.cin);
        w.raw(l.weights.data(), 2 * l.weights.size()); w.raw(l.bias.data(), 2 * l.bias.size());
        w.str(kActivationNames[l.activation]); w.put<float>(l.activation_param);
    }
    w.raw(box_min, 12); w.raw(box_size, 12);
    w.put<uint8_t>(has_latent_grid ? 1 : 0);
    if (has_latent_grid) {
        w.put<int32_t>(1);
        w.put<int32_t>(time_min); w.put<int32_t>(time_num); w.put<int32_t>(time_step); w.put<int32_t>(ensemble_min); w.put<int32_t>(ensemble_num);
        for (const auto& g : time_grids) save_grid(w, g);
        for (const auto& g : ensemble_grids) save_grid(w, g);
    }
    return w.buf;
}

} // namespace fvsrn
