#include "volume.hpp"
#include <cstring>
#include <stdexcept>

namespace fvsrn {
namespace {

struct Reader {
    const uint8_t* p; size_t n, pos = 0;
    void need(size_t k) const { if (pos + k > n) throw std::runtime_error("cvol: unexpected end of data"); }
    template <class T> T get() { need(sizeof(T)); T v; std::memcpy(&v, p + pos, sizeof(T)); pos += sizeof(T); return v; }
    void skip(size_t k) { need(k); pos += k; }
    const uint8_t* take(size_t k) { need(k); const uint8_t* r = p + pos; pos += k; return r; }
};

template <class T> void put(std::vector<uint8_t>& o, T v) { const uint8_t* b = reinterpret_cast<const uint8_t*>(&v); o.insert(o.end(), b, b + sizeof(T)); }

// body of a feature: raw z-slices, or a sequence of [int32 compressed size][LZ4 block] chunks (lz4cpp.cpp:94-102)
void read_body(Reader& r, std::vector<uint8_t>& dst, bool compressed) {
    if (!compressed) { std::memcpy(dst.data(), r.take(dst.size()), dst.size()); return; }
    size_t off = 0;
    while (off < dst.size()) {
        const int32_t len = r.get<int32_t>();
        if (len <= 0) throw std::runtime_error("cvol: bad compressed chunk size");
        const size_t got = lz4_decode_block(r.take(size_t(len)), size_t(len), dst.data(), off, dst.size());
        if (got == 0) throw std::runtime_error("cvol: empty compressed chunk");
        off += got;
    }
}

} // namespace

size_t lz4_decode_block(const uint8_t* src, size_t src_len, uint8_t* dst, size_t dst_pos, size_t dst_cap) {
    size_t ip = 0, op = dst_pos;
    while (ip < src_len) {
        const unsigned token = src[ip++];
        size_t lit = token >> 4;
        if (lit == 15) { unsigned b; do { if (ip >= src_len) throw std::runtime_error("lz4: truncated literal length"); b = src[ip++]; lit += b; } while (b == 255); }
        if (ip + lit > src_len || op + lit > dst_cap) throw std::runtime_error("lz4: literal run out of bounds");
        std::memcpy(dst + op, src + ip, lit); ip += lit; op += lit;
        if (ip >= src_len) break;                                   // the last sequence has no match part
        if (ip + 2 > src_len) throw std::runtime_error("lz4: truncated offset");
        const size_t offset = size_t(src[ip]) | (size_t(src[ip + 1]) << 8); ip += 2;
        if (offset == 0 || offset > op) throw std::runtime_error("lz4: bad match offset");
        size_t mlen = token & 15;
        if (mlen == 15) { unsigned b; do { if (ip >= src_len) throw std::runtime_error("lz4: truncated match length"); b = src[ip++]; mlen += b; } while (b == 255); }
        mlen += 4;
        if (op + mlen > dst_cap) throw std::runtime_error("lz4: match out of bounds");
        for (size_t k = 0; k < mlen; ++k, ++op) dst[op] = dst[op - offset];    // byte-wise: matches may overlap their own output
    }
    return op - dst_pos;
}

float VolumeFeature::value(uint64_t x, uint64_t y, uint64_t z, int c) const {
    const size_t i = size_t(c) + size_t(channels) * (x + X * (y + Y * z));
    if (type == VOL_UCHAR) return float(data[i]) / 255.0f;
    if (type == VOL_USHORT) { uint16_t v; std::memcpy(&v, data.data() + 2 * i, 2); return float(v) / 65535.0f; }
    float v; std::memcpy(&v, data.data() + 4 * i, 4); return v;
}

Volume Volume::load(const uint8_t* bytes, size_t n) {
    Reader r{bytes, n};
    Volume v;
    const uint8_t* magic = r.take(4);
    if (std::memcmp(magic, "CVOL", 4) == 0) {                           // volume.cpp:710-740
        if (r.get<int32_t>() != 1) throw std::runtime_error("Unknown file version!");
        for (int i = 0; i < 3; ++i) v.world[i] = r.get<float>();
        const int32_t nf = r.get<int32_t>(), flags = r.get<int32_t>();
        r.skip(4);
        if (nf < 0 || nf > 4096) throw std::runtime_error("cvol: bad feature count");
        for (int f = 0; f < nf; ++f) {                                   // Feature::load, :278-332
            VolumeFeature ft;
            const int32_t len = r.get<int32_t>();
            if (len < 0 || len > 4096) throw std::runtime_error("cvol: bad feature name");
            ft.name.assign(reinterpret_cast<const char*>(r.take(size_t(len))), size_t(len));
            ft.X = r.get<uint64_t>(); ft.Y = r.get<uint64_t>(); ft.Z = r.get<uint64_t>();
            ft.channels = r.get<int32_t>(); ft.type = r.get<int32_t>();
            if (ft.type < 0 || ft.type > VOL_FLOAT || ft.channels < 1 || ft.channels > 4 || ft.X * ft.Y * ft.Z == 0 || ft.X * ft.Y * ft.Z > (1ull << 33))
                throw std::runtime_error("cvol: bad feature header");
            ft.data.resize(ft.bytes_per_entry() * ft.X * ft.Y * ft.Z * size_t(ft.channels));
            read_body(r, ft.data, (flags & 1) != 0);
            v.features.push_back(std::move(ft));
        }
    } else if (std::memcmp(magic, "cvol", 4) == 0) {                    // legacy container: one density feature, :741-790
        VolumeFeature ft; ft.name = "density";
        ft.X = r.get<uint64_t>(); ft.Y = r.get<uint64_t>(); ft.Z = r.get<uint64_t>();
        const double vx = r.get<double>(), vy = r.get<double>(), vz = r.get<double>();
        ft.type = int(r.get<uint32_t>());
        const bool compressed = r.get<uint8_t>() != 0;
        r.skip(7);
        if (ft.type < 0 || ft.type > VOL_FLOAT || ft.X * ft.Y * ft.Z == 0 || ft.X * ft.Y * ft.Z > (1ull << 33)) throw std::runtime_error("cvol: bad header");
        v.world[0] = float(vx * double(ft.X)); v.world[1] = float(vy * double(ft.Y)); v.world[2] = float(vz * double(ft.Z));
        ft.data.resize(ft.bytes_per_entry() * ft.X * ft.Y * ft.Z);
        read_body(r, ft.data, compressed);
        v.features.push_back(std::move(ft));
    } else throw std::runtime_error("Illegal magic number");
    return v;
}

std::vector<uint8_t> Volume::save() const {                              // Volume::save, :622-670, compression 0
    std::vector<uint8_t> o;
    o.insert(o.end(), {'C', 'V', 'O', 'L'});
    put<int32_t>(o, 1);
    for (int i = 0; i < 3; ++i) put<float>(o, world[i]);
    put<int32_t>(o, int32_t(features.size())); put<int32_t>(o, 0); put<int32_t>(o, 0);
    for (const auto& f : features) {                                     // Feature::save, :334-384
        put<int32_t>(o, int32_t(f.name.size()));
        o.insert(o.end(), f.name.begin(), f.name.end());
        put<uint64_t>(o, f.X); put<uint64_t>(o, f.Y); put<uint64_t>(o, f.Z);
        put<int32_t>(o, f.channels); put<int32_t>(o, f.type);
        o.insert(o.end(), f.data.begin(), f.data.end());
    }
    return o;
}

Volume Volume::from_dense(const float* data, int X, int Y, int Z, const float world_size[3], const char* name) {
    if (!data || X <= 0 || Y <= 0 || Z <= 0) throw std::runtime_error("volume: bad dense grid");
    Volume v;
    for (int i = 0; i < 3; ++i) v.world[i] = world_size[i];
    VolumeFeature f; f.name = name; f.type = VOL_FLOAT; f.channels = 1; f.X = uint64_t(X); f.Y = uint64_t(Y); f.Z = uint64_t(Z);
    f.data.resize(size_t(X) * Y * Z * 4);
    std::memcpy(f.data.data(), data, f.data.size());
    v.features.push_back(std::move(f));
    return v;
}

} // namespace fvsrn
