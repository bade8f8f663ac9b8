// Host side of the ground-truth volume: the .cvol container of the reference (renderer/volume.h:42-77,
// volume.cpp:278-384, 618-800) - what the eval scripts render their comparison images from. Written from the format
// description in volume.h; the LZ4 block decoder follows the published LZ4 block format (the reference links liblz4).
#pragma once
#include <cstdint>
#include <string>
#include <vector>

namespace fvsrn {

enum { VOL_UCHAR = 0, VOL_USHORT = 1, VOL_FLOAT = 2 };      // Volume::DataType, volume.h:83-89

struct VolumeFeature {
    std::string name;
    int type = VOL_FLOAT;
    int channels = 1;
    uint64_t X = 0, Y = 0, Z = 0;
    std::vector<uint8_t> data;      // index = channel + channels * (x + X * (y + Y * z)), volume.h:126-132
    size_t bytes_per_entry() const { return type == VOL_UCHAR ? 1 : (type == VOL_USHORT ? 2 : 4); }
    // value as the texture unit returns it: floats as stored, integer types normalised to [0,1] (volume.cpp:155)
    float value(uint64_t x, uint64_t y, uint64_t z, int c = 0) const;
};

struct Volume {
    float world[3] = {1.f, 1.f, 1.f};       // size of the bounding box in world space
    std::vector<VolumeFeature> features;

    static Volume load(const uint8_t* bytes, size_t n);     // "CVOL" (version 1) and the legacy "cvol" container, LZ4 or raw
    std::vector<uint8_t> save() const;                       // "CVOL" version 1, uncompressed
    static Volume from_dense(const float* data, int X, int Y, int Z, const float world_size[3], const char* name = "density");
};

// LZ4 block -> dst (which may already hold `dst_pos` bytes of earlier blocks: matches reach back into them, the
// reference decodes with LZ4_decompress_safe_continue into one contiguous buffer). Returns the bytes written.
size_t lz4_decode_block(const uint8_t* src, size_t src_len, uint8_t* dst, size_t dst_pos, size_t dst_cap);

} // namespace fvsrn
