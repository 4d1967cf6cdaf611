"""4c_b200 — B200-native evaluation + assembly of 4C's 3D solid elements (hex8 / hex27 / tet10,
displacement-based and F-bar, St.Venant-Kirchhoff and compressible Neo-Hooke).

The product is the CUDA library `lib/libfourc_b200.so` (sources in `csrc/`, C ABI in
`include/fourc_b200.h`).  This package is its Python plumbing: `capi` (ctypes binding: `Solid`, `Grid`).
The C++ mirror of the reference's Discretization calling sequence is `host/fourc_b200.hpp`.  The directory
name starts with a digit, so import it with `importlib.import_module("4c_b200")`.
"""
from . import capi  # noqa: F401
from .capi import *  # noqa: F401,F403
