"""
The ground-truth volume container: .cvol reading / writing through the C ABI (renderer/volume.cpp:278-384,
618-800; format renderer/volume.h:42-77), CPU only. The LZ4-compressed bodies are decoded by the library's own block
decoder; the test compresses with a small independent LZ4 encoder written from the block-format description.
"""
import ctypes as C
import os
import struct

import numpy as np
import pytest

from fvsrn_b200 import _lib

REAL = "/root/reference/applications/volumes/RichtmyerMeshkov/ppm-t0020.cvol"


def lz4_block(data: bytes) -> bytes:
    """Greedy LZ4 block encoder (4-byte hash-free matcher over a small window): enough to exercise literals, short and long
    (>= 19) matches, overlapping matches and the end-of-block rules (last 5 bytes literals, last match 12 bytes before the end)."""
    out = bytearray(); i = 0; anchor = 0; n = len(data)

    def emit(lit, mlen, off):
        token_l = min(len(lit), 15); token_m = min(mlen - 4, 15) if mlen else 0
        out.append((token_l << 4) | token_m)
        if len(lit) >= 15:
            r = len(lit) - 15
            while r >= 255: out.append(255); r -= 255
            out.append(r)
        out.extend(lit)
        if mlen:
            out.extend(struct.pack("<H", off))
            if mlen - 4 >= 15:
                r = mlen - 4 - 15
                while r >= 255: out.append(255); r -= 255
                out.append(r)
    while i + 12 < n:
        best = 0; boff = 0
        for off in range(1, min(i, 300) + 1):
            l = 0
            while i + l < n - 5 and data[i + l] == data[i + l - off] and l < 400: l += 1
            if l > best: best, boff = l, off
        if best >= 4:
            emit(data[anchor:i], best, boff); i += best; anchor = i
        else:
            i += 1
    emit(data[anchor:], 0, 0)
    return bytes(out)


def chunks(data: bytes, size: int) -> bytes:
    """LZ4Compressor framing (third-party/lz4/lib/lz4cpp.cpp): [int32 compressed size][block] per chunk."""
    out = b""
    for o in range(0, len(data), size):
        b = lz4_block(data[o:o + size])
        out += struct.pack("<i", len(b)) + b
    return out


def load(L, data):
    h = C.c_void_p()
    _lib.check(L.fvsrn_volume_load_cvol(data, len(data), C.byref(h)))
    return h


def save(L, h):
    n = C.c_size_t()
    _lib.check(L.fvsrn_volume_save_cvol(h, None, 0, C.byref(n)))
    buf = C.create_string_buffer(n.value)
    _lib.check(L.fvsrn_volume_save_cvol(h, buf, n.value, C.byref(n)))
    return buf.raw[:n.value]


def test_cvol_roundtrip_dense_and_formats(built_lib):
    L = built_lib
    rng = np.random.default_rng(0)
    vol = rng.random((5, 6, 7), dtype=np.float32)                   # (Z,Y,X)
    h = C.c_void_p()
    _lib.check(L.fvsrn_volume_from_dense(vol.ctypes.data, 7, 6, 5, C.byref((C.c_float * 3)(1.0, 2.0, 0.5)), C.byref(h)))
    info = _lib.VolumeInfo(); _lib.check(L.fvsrn_volume_get_info(h, 0, C.byref(info)))
    assert (info.size_x, info.size_y, info.size_z, info.channels, info.data_type, info.num_features) == (7, 6, 5, 1, 2, 1)
    assert list(info.box_min) == [-0.5, -1.0, -0.25] and list(info.box_size) == [1.0, 2.0, 0.5]
    data = save(L, h)
    # header per renderer/volume.h:44-61
    assert data[:4] == b"CVOL" and struct.unpack("<i3fiii", data[4:32]) == (1, 1.0, 2.0, 0.5, 1, 0, 0)
    name_len = struct.unpack("<i", data[32:36])[0]
    assert data[36:36 + name_len] == b"density"
    body = data[36 + name_len + 32:]
    assert np.array_equal(np.frombuffer(body, np.float32).reshape(5, 6, 7), vol)
    h2 = load(L, data)
    assert save(L, h2) == data
    L.fvsrn_volume_free(h); L.fvsrn_volume_free(h2)
    assert L.fvsrn_volume_get_info(None, 0, C.byref(info)) != 0
    with pytest.raises(RuntimeError, match="Illegal magic number"):
        load(L, b"nope" + bytes(60))
    with pytest.raises(RuntimeError, match="unexpected end"):
        load(L, data[:100])


@pytest.mark.parametrize("legacy", [False, True])
def test_cvol_lz4_compressed(built_lib, legacy):
    L = built_lib
    rng = np.random.default_rng(1)
    # compressible uchar volume: smooth ramps + runs (long overlapping matches) + noise (literals)
    z, y, x = np.meshgrid(np.arange(9), np.arange(10), np.arange(33), indexing="ij")
    vol = ((x * 3 + y * 7 + z * 11) % 64).astype(np.uint8)
    vol[4:] = 17
    vol[2, :, :16] = rng.integers(0, 255, (10, 16), dtype=np.uint8)
    raw = vol.tobytes()
    body = chunks(raw, 1000)                                        # several chunks: later blocks reference earlier output
    assert len(body) < len(raw)
    if legacy:      # "cvol": sizes u64 x3, voxel sizes f64 x3, type u32, compression flag, 7 pad bytes (volume.cpp:741-757)
        data = b"cvol" + struct.pack("<QQQdddIB7x", 33, 10, 9, 1 / 33, 1 / 10, 1 / 9, 0, 1) + body
    else:
        data = b"CVOL" + struct.pack("<i3fiii", 1, 1.0, 1.0, 1.0, 1, 1, 0) + struct.pack("<i", 3) + b"rho" + struct.pack("<QQQii", 33, 10, 9, 1, 0) + body
    h = load(L, data)
    info = _lib.VolumeInfo(); _lib.check(L.fvsrn_volume_get_info(h, 0, C.byref(info)))
    assert (info.size_x, info.size_y, info.size_z, info.data_type) == (33, 10, 9, 0)
    assert np.allclose(list(info.world_size), 1.0)
    out = save(L, h)
    assert out[-len(raw):] == raw                                   # saved uncompressed: the decoded voxels
    L.fvsrn_volume_free(h)


@pytest.mark.skipif(not os.path.exists(REAL), reason="the reference tree is only present in the build container")
def test_reads_the_reference_example_volume(built_lib):
    """applications/volumes/RichtmyerMeshkov/ppm-t0020.cvol: legacy container, 256^3 uchar, LZ4."""
    L = built_lib
    h = load(L, open(REAL, "rb").read())
    info = _lib.VolumeInfo(); _lib.check(L.fvsrn_volume_get_info(h, 0, C.byref(info)))
    assert (info.size_x, info.size_y, info.size_z, info.channels, info.data_type) == (256, 256, 256, 1, 0)
    assert np.allclose(list(info.world_size), 1.0) and np.allclose(list(info.box_min), -0.5)
    vox = np.frombuffer(save(L, h)[-256 ** 3:], np.uint8)
    assert vox.min() < vox.max() and 0 < vox.mean() < 255
    L.fvsrn_volume_free(h)
