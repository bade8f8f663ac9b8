"""
fvsrn_b200 — B200-native fused SRN evaluation + DVR ray-marching (the hot path of shamanDevel/fV-SRN).

The directory is named ``fv-srn_b200`` (not an identifier); import it as ``fvsrn_b200`` through the loader shim
``fvsrn_b200.py`` at the repository root.

Sub-modules: ``pyrenderer`` (host-side mirror of the reference's pybind11 module, compute through the C ABI),
``volnet`` (.volnet reader/writer), ``scenes`` (synthetic benchmark scenes), ``_lib`` (ctypes binding).
"""
from . import _lib, volnet  # noqa: F401

__all__ = ["_lib", "volnet", "pyrenderer", "scenes"]


def __getattr__(name):  # torch is imported lazily: volnet / ABI checks do not need it
    if name in ("pyrenderer", "scenes", "dist"):
        import importlib
        return importlib.import_module(f"{__name__}.{name}")
    raise AttributeError(name)
