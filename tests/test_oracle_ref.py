"""The oracle's dense restatement, bit for bit against the REFERENCE's own header-only kernels
compiled from /root/reference into oracle/_ref/libfc_ref_dense.so (oracle/ref_dense_harness.cpp).
The prebuilt .so travels to the GPU box; the sources never enter this repo."""
import ctypes as C
import os

import numpy as np
import oracle_lib as O
import pytest

REF = os.path.join(O.ORACLE_DIR, "_ref", "libfc_ref_dense.so")
pytestmark = pytest.mark.skipif(not os.path.exists(REF), reason="oracle/_ref not built (no /root/reference here)")


def p(a):
    return a.ctypes.data_as(C.c_void_p)


@pytest.fixture(scope="module")
def ref():
    r = C.CDLL(REF)
    for n in ("ref_invert3x3", "ref_determinant3x3", "ref_voigt_determinant"):
        getattr(r, n).restype = C.c_double
    for n in ("oracle_invert3x3", "oracle_determinant3x3", "oracle_voigt_determinant"):
        getattr(O.lib, n).restype = C.c_double
    return r


def test_multiply_bit_exact(ref):
    rng = np.random.default_rng(1)
    for _ in range(20):
        c, B = rng.standard_normal(36), rng.standard_normal(6 * 24)
        a, b = np.zeros(6 * 24), np.zeros(6 * 24)
        ref.ref_multiply_nn_6_6_24(p(a), p(c), p(B))
        O.lib.oracle_multiply_nn_6_6_24(p(b), p(c), p(B))
        assert np.array_equal(a, b)
        K0 = rng.standard_normal(576)
        Ka, Kb = K0.copy(), K0.copy()
        ref.ref_multiply_tn_24_6_24(C.c_double(0.37), p(Ka), p(B), p(a))
        O.lib.oracle_multiply_tn_24_6_24(C.c_double(0.37), p(Kb), p(B), p(a))
        assert np.array_equal(Ka, Kb)
        dN, X = rng.standard_normal(24), rng.standard_normal(24)
        Ja, Jb = np.zeros(9), np.zeros(9)
        ref.ref_multiply_nt_3_8_3(p(Ja), p(dN), p(X))
        O.lib.oracle_multiply_nt_3_8_3(p(Jb), p(dN), p(X))
        assert np.array_equal(Ja, Jb)
        F = rng.standard_normal(9)
        Ca, Cb = np.zeros(9), np.zeros(9)
        ref.ref_multiply_tn_3_3_3(p(Ca), p(F), p(F))
        O.lib.oracle_multiply_tn_3_3_3(p(Cb), p(F), p(F))
        assert np.array_equal(Ca, Cb)


def test_invert_determinant_voigt_bit_exact(ref):
    rng = np.random.default_rng(2)
    for _ in range(50):
        M = np.eye(3).reshape(-1) + 0.3 * rng.standard_normal(9)
        ia, ib = np.zeros(9), np.zeros(9)
        da = ref.ref_invert3x3(p(M), p(ia))
        db = O.lib.oracle_invert3x3(p(M), p(ib))
        assert da == db and np.array_equal(ia, ib)
        assert ref.ref_determinant3x3(p(M)) == O.lib.oracle_determinant3x3(p(M))
        t = np.array([1.1, 1.2, 1.3, 0.0, 0.0, 0.0]) + 0.1 * rng.standard_normal(6)
        for strain in (0, 1):
            assert ref.ref_voigt_determinant(C.c_int(strain), p(t)) == O.lib.oracle_voigt_determinant(C.c_int(strain), p(t))
            pa, pb = np.zeros(3), np.zeros(3)
            ref.ref_voigt_invariants(C.c_int(strain), p(t), p(pa))
            O.lib.oracle_voigt_invariants(C.c_int(strain), p(t), p(pb))
            assert np.array_equal(pa, pb)
