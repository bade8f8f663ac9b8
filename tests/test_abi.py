"""C-ABI checks that need no GPU: the library loads, exports every symbol include/fvsrn_b200.h declares, the host-side
.volnet model behaves like renderer::SceneNetwork, and compute entry points fail loudly without a CUDA device."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from conftest import ROOT, Golden, golden_names

from fvsrn_b200 import _lib, scenes, volnet


def _declared_symbols():
    src = open(os.path.join(ROOT, "include", "fvsrn_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(fvsrn_[a-z_0-9]+)\s*\(", src)))


def test_header_and_binding_agree():
    assert _declared_symbols() == sorted(_lib.SYMBOLS)


def test_library_exports_every_declared_symbol(built_lib):
    for s in _declared_symbols():
        assert hasattr(built_lib, s), s
    assert built_lib.fvsrn_abi_version() == 3


def test_struct_sizes_match_header(built_lib):
    # natural C layout of the header structs on LP64 ...
    assert C.sizeof(_lib.NetworkInfo) == 22 * 4
    assert C.sizeof(_lib.DvrParams) % 8 == 0 and C.sizeof(_lib.DvrParams) == 216
    assert C.sizeof(_lib.QueryInfo) == 48
    # ... and as the C compiler lays out include/fvsrn_b200.h
    import subprocess, tempfile
    with tempfile.TemporaryDirectory() as d:
        src = os.path.join(d, "s.c")
        open(src, "w").write('#include <stdio.h>\n#include <stddef.h>\n#include "fvsrn_b200.h"\nint main(void){printf("%zu %zu %zu %zu %zu\\n", '
                             'sizeof(fvsrn_network_info), sizeof(fvsrn_dvr_params), sizeof(fvsrn_query_info), '
                             'offsetof(fvsrn_dvr_params, output_format), offsetof(fvsrn_dvr_params, out_stride_b));return 0;}\n')
        subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), src, "-o", os.path.join(d, "s")])
        sizes = [int(v) for v in subprocess.check_output([os.path.join(d, "s")]).split()]
    assert sizes == [C.sizeof(_lib.NetworkInfo), C.sizeof(_lib.DvrParams), C.sizeof(_lib.QueryInfo),
                     _lib.DvrParams.output_format.offset, _lib.DvrParams.out_stride_b.offset]


def _load(built_lib, data):
    h = C.c_void_p()
    _lib.check(built_lib.fvsrn_network_load_volnet(data, len(data), C.byref(h)))
    return h


def _save(built_lib, h):
    n = C.c_size_t()
    _lib.check(built_lib.fvsrn_network_save_volnet(h, None, 0, C.byref(n)))
    buf = C.create_string_buffer(n.value)
    _lib.check(built_lib.fvsrn_network_save_volnet(h, buf, n.value, C.byref(n)))
    return buf.raw[:n.value]


@pytest.mark.parametrize("name", golden_names())
def test_volnet_roundtrip_through_c_abi(built_lib, name):
    """testSRN.cpp:413-430 through the C ABI: load -> save is byte-identical; info matches the Python parser."""
    g = Golden(name)
    h = _load(built_lib, g.volnet_bytes)
    try:
        assert _save(built_lib, h) == g.volnet_bytes
        info = _lib.NetworkInfo()
        _lib.check(built_lib.fvsrn_network_get_info(h, C.byref(info)))
        assert info.hidden_channels == g.net.input_channels_out()
        assert info.num_layers == len(g.net.layers)
        assert info.latent_channels == g.net.latent_channels()
        assert info.num_fourier == g.net.num_fourier
        assert info.channels_out == g.world.shape[1]
        assert built_lib.fvsrn_network_valid(h) == 1
    finally:
        built_lib.fvsrn_network_free(h)


def test_builders_match_python_export(built_lib):
    """Building through fvsrn_network_add_layer & co. gives the same bytes as the Python writer (addLayer re-layout,
    volume_interpolation_network.cpp:806-894: zero column at input 3, transposed last layer)."""
    rng = np.random.default_rng(5)
    H, nf, gc, R = 32, 14, 16, 4
    B = rng.normal(0, 1, (nf, 3)).astype(np.float32)
    Ws = [rng.normal(0, .3, (H, 3 + 2 * nf + gc)).astype(np.float32), rng.normal(0, .3, (H, H)).astype(np.float32),
          rng.normal(0, .3, (1, H)).astype(np.float32)]
    bs = [rng.normal(0, .3, (w.shape[0],)).astype(np.float32) for w in Ws]
    grid = rng.normal(0, 1, (gc, R, R, R)).astype(np.float32)
    # python
    pn = volnet.SceneNetwork()
    pn.set_fourier_matrix(B, premultiplied=False)
    pn.output_mode = "density"
    for i, (w, b) in enumerate(zip(Ws, bs)):
        pn.add_layer(w, b, "SnakeAlt" if i < 2 else "None", 2.0 if i < 2 else 1.0)
    pn.has_latent_grid = True
    pn.time_min, pn.time_num, pn.time_step, pn.ensemble_min, pn.ensemble_num = 0, 1, 1, 0, 0
    pn.time_grids = [volnet.LatentGrid(volnet.GRID_FLOAT, grid)]
    pn.box_min, pn.box_size = (-0.5, -0.5, -0.5), (1.0, 1.0, 1.0)
    assert pn.valid()
    # C ABI
    h = C.c_void_p()
    L = built_lib
    _lib.check(L.fvsrn_network_create(C.byref(h)))
    try:
        assert L.fvsrn_network_valid(h) == 0
        _lib.check(L.fvsrn_network_set_input(h, 0, 0))
        _lib.check(L.fvsrn_network_set_fourier(h, B.ctypes.data, nf, 3, 0))
        _lib.check(L.fvsrn_network_set_output_mode(h, b"density"))
        for i, (w, b) in enumerate(zip(Ws, bs)):
            _lib.check(L.fvsrn_network_add_layer(h, w.ctypes.data, b.ctypes.data, w.shape[0], w.shape[1],
                                                 b"SnakeAlt" if i < 2 else b"None", 2.0 if i < 2 else 1.0))
        _lib.check(L.fvsrn_network_set_latent_grid_layout(h, 0, 1, 1, 0, 0))
        err = C.c_double(-1)
        _lib.check(L.fvsrn_network_set_time_grid(h, 0, grid.ctypes.data, gc, R, R, R, 0, C.byref(err)))
        assert err.value == 0.0
        _lib.check(L.fvsrn_network_set_box(h, C.byref((C.c_float * 3)(-.5, -.5, -.5)), C.byref((C.c_float * 3)(1, 1, 1))))
        assert L.fvsrn_network_valid(h) == 1
        assert _save(L, h) == volnet.dumps(pn)
        assert L.fvsrn_network_num_parameters(h) > 0
        assert 1 <= L.fvsrn_network_compute_max_warps(h, 0, 0)
    finally:
        L.fvsrn_network_free(h)


def test_errors_surface_as_runtime_error(built_lib):
    L = built_lib
    h = C.c_void_p()
    with pytest.raises(RuntimeError):
        _lib.check(L.fvsrn_network_load_volnet(b"\x01\x00\x00\x00garbage", 11, C.byref(h)))
    _lib.check(L.fvsrn_network_create(C.byref(h)))
    try:
        with pytest.raises(RuntimeError, match="output mode"):
            _lib.check(L.fvsrn_network_set_output_mode(h, b"nonsense"))
        w = np.zeros((4, 4), np.float32)
        with pytest.raises(RuntimeError):
            _lib.check(L.fvsrn_network_add_layer(h, w.ctypes.data, w.ctypes.data, 4, 4, b"NoSuchActivation", 1.0))
        B = np.zeros((4, 5), np.float32)
        with pytest.raises(RuntimeError, match="expected: 3 or 6"):
            _lib.check(L.fvsrn_network_set_fourier(h, B.ctypes.data, 4, 5, 1))
    finally:
        L.fvsrn_network_free(h)


def test_no_cpu_fallback(built_lib):
    """Without a CUDA device prepare() must fail loudly; evaluate/render on an unprepared handle must fail."""
    import torch
    g = Golden("c1_relu_nerf")
    h = _load(built_lib, g.volnet_bytes)
    try:
        out = np.zeros(4, np.float32)
        with pytest.raises(RuntimeError, match="not prepared"):
            _lib.check(built_lib.fvsrn_evaluate(h, out.ctypes.data, None, out.ctypes.data, 1, None))
        with pytest.raises(RuntimeError, match="not prepared"):
            _lib.check(built_lib.fvsrn_evaluate_host(h, out.ctypes.data, None, out.ctypes.data, 1, None))
        if not torch.cuda.is_available():
            with pytest.raises(RuntimeError, match="no CPU fallback"):
                _lib.check(built_lib.fvsrn_network_prepare(h, 0, None))
    finally:
        built_lib.fvsrn_network_free(h)


def test_product_package_never_touches_the_oracle():
    pkg = os.path.join(ROOT, "fv-srn_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".hpp", ".h")):
                text = open(os.path.join(dirpath, f), errors="replace").read()
                assert "import oracle" not in text and "fvsrn_oracle" not in text, f
