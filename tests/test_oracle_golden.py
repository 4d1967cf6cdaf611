"""The CPU oracle against the reference's own golden vectors / known-answer tests.
Each test names the reference test it restates; values are the literals of those files."""
import ctypes as C

import numpy as np
import oracle_lib as O

L = O.lib


def p(a):
    return a.ctypes.data_as(C.c_void_p)


def test_calculate_gamma_delta():
    # unittests/mat/4C_elasthyper_service_test.cpp:36-82 (prinv_ = 3.1, 3.2, 1.05 in the fixture)
    g, d = O.gamma_delta([3.1, 3.2, 1.05], [0.536, 0.681, 3.981], [0.153, 0.861, 2.964, 1.549, 0.682, 3.584])
    np.testing.assert_allclose(g, [5.294200000000001, -1.362, 8.360099999999999], atol=1e-10, rtol=0)
    np.testing.assert_allclose(d, [125.31604, -25.0124, 23.03238, 3.444, -6.5058, 29.79144, -16.7202, -2.724], atol=1e-10, rtol=0)


def test_right_cauchy_green_strain_like():
    # unittests/mat/4C_elasthyper_service_test.cpp:84-109
    E = np.array([0.1, 0.2, 0.3, 0.01, 0.02, 0.03])
    Cv = np.zeros(6)
    L.oracle_rcg_strain_like(p(E), p(Cv))
    np.testing.assert_allclose(Cv, [1.2, 1.4, 1.6, 0.02, 0.04, 0.06], atol=1e-10, rtol=0)


def test_stvenantkirchhoff_stress_and_energy():
    # unittests/mat/4C_stvenantkirchhoff_test.cpp:41-63, 80-128: E=210, nu=0.3, glstrain = ones
    young, nu = 210.0, 0.3
    S, cmat = O.material_evaluate(O.MAT_SVK, young, nu, None, np.ones(6))
    n = young / ((1 + nu) * (1 - 2 * nu)) * ((1 - nu) + nu + nu)
    s = young / ((1 + nu) * (1 - 2 * nu)) * ((1 - 2 * nu) / 2)
    np.testing.assert_allclose(S, [n, n, n, s, s, s], atol=1e-4, rtol=0)
    L.oracle_svk_strain_energy.restype = C.c_double
    psi = L.oracle_svk_strain_energy(C.c_double(young), C.c_double(nu), p(np.ones(6)))
    assert abs(psi - 908.6538) < 1e-4


TENS_STRESS = np.array([1.1, 1.2, 1.3, 0.01, 0.02, 0.03])
TENS_STRAIN = np.array([1.1, 1.2, 1.3, 0.02, 0.04, 0.06])
ITENS_STRESS = np.array([0.90972618, 0.83360457, 0.76990741, -0.00723301, -0.01265777, -0.0208824])


def test_voigt_determinant_invariants_inverse():
    # src/core/linalg/tests/4C_linalg_utils_voigt_notation_test.cpp:72-131
    L.oracle_voigt_determinant.restype = C.c_double
    for strain, t in ((0, TENS_STRESS), (1, TENS_STRAIN)):
        assert abs(L.oracle_voigt_determinant(C.c_int(strain), p(t)) - 1.7143620000000002) < 1e-10
        pr = np.zeros(3)
        L.oracle_voigt_invariants(C.c_int(strain), p(t), p(pr))
        np.testing.assert_allclose(pr, [3.5999999999999996, 4.3085999999999984, 1.7143620000000002], atol=1e-10, rtol=0)
        ti = np.zeros(6)
        L.oracle_voigt_inverse(C.c_int(strain), p(t), p(ti))
        ref = ITENS_STRESS * (np.array([1, 1, 1, 2, 2, 2]) if strain else 1)
        np.testing.assert_allclose(ti, ref, atol=1e-5, rtol=0)
    out = np.zeros(6)
    L.oracle_voigt_to_stress_like(p(TENS_STRAIN), p(out))
    np.testing.assert_allclose(out, TENS_STRESS, atol=1e-10, rtol=0)
    m = np.array([[1.1, 0.01, 0.03], [0.01, 1.2, 0.02], [0.03, 0.02, 1.3]])
    v = np.zeros(6)
    L.oracle_voigt_matrix_to_vector_strain(p(np.ascontiguousarray(m.T.reshape(-1))), p(v))
    np.testing.assert_allclose(v, TENS_STRAIN, atol=1e-10, rtol=0)


def test_holzapfel_product_index_rule():
    # no reference unit test exists for add_holzapfel_product
    # (src/core/linalg/src/dense/4C_linalg_fixedsizematrix_tensor_products.cpp:264-314): checked against its
    # defining rule (A (.) A)_ijkl = 1/2 (A_ik A_jl + A_il A_jk) in stress/stress Voigt notation
    A = np.array([[0.9, -0.007, -0.02], [-0.007, 0.83, -0.012], [-0.02, -0.012, 0.77]])
    v = np.array([A[0, 0], A[1, 1], A[2, 2], A[0, 1], A[1, 2], A[0, 2]])
    cm = np.zeros(36)
    L.oracle_add_holzapfel_product(p(cm), p(v), C.c_double(1.7))
    cm = cm.reshape(6, 6).T
    pairs = [(0, 0), (1, 1), (2, 2), (0, 1), (1, 2), (0, 2)]
    ref = np.array([[1.7 * 0.5 * (A[i, k] * A[j, l] + A[i, l] * A[j, k]) for (k, l) in pairs] for (i, j) in pairs])
    np.testing.assert_allclose(cm, ref, atol=1e-15, rtol=1e-14)


def test_coupneohooke_against_closed_form():
    # MAT_ElastHyper + ELAST_CoupNeoHooke (src/mat/elast/4C_mat_elast_coupneohooke.cpp:60-77):
    # S = 2c (I - I3^-beta C^-1), checked on the material of tests/input_files/solid_ele_hex8_Standard_eas_none_dynamic.dat
    young, nu = 10.0, 0.25
    c, beta = young / (4 * (1 + nu)), nu / (1 - 2 * nu)
    F = np.array([[1.1, 0.05, 0.0], [0.02, 0.95, 0.03], [0.0, 0.01, 1.05]])
    Cm = F.T @ F
    Ev = np.array([0.5 * (Cm[0, 0] - 1), 0.5 * (Cm[1, 1] - 1), 0.5 * (Cm[2, 2] - 1), Cm[0, 1], Cm[1, 2], Cm[0, 2]])
    S, cmat = O.material_evaluate(O.MAT_NEOHOOKE, young, nu, F, Ev)
    Sm = 2 * c * (np.eye(3) - np.linalg.det(Cm) ** (-beta) * np.linalg.inv(Cm))
    np.testing.assert_allclose(S, [Sm[0, 0], Sm[1, 1], Sm[2, 2], Sm[0, 1], Sm[1, 2], Sm[0, 2]], atol=1e-13)
    # tangent = 2 dS/dC by central differences on the strain-like Voigt vector
    h = 1e-6
    for k in range(6):
        dE = np.zeros(6)
        dE[k] = h
        Sp, _ = O.material_evaluate(O.MAT_NEOHOOKE, young, nu, F, Ev + dE)
        Sm_, _ = O.material_evaluate(O.MAT_NEOHOOKE, young, nu, F, Ev - dE)
        np.testing.assert_allclose(cmat[:, k], (Sp - Sm_) / (2 * h), atol=2e-8)


def test_gauss_rules_and_partition_of_unity():
    for ct, vol in ((O.HEX8, 8.0), (O.HEX27, 8.0), (O.TET10, 1.0 / 6.0), (O.HEX20, 8.0), (O.TET4, 1.0 / 6.0), (O.WEDGE6, 1.0), (O.PYRAMID5, 4.0 / 3.0)):
        n, g = O.NEN[ct], O.NGP[ct]
        xi, w = np.zeros(3 * g), np.zeros(g)
        L.oracle_gauss_rule(C.c_int(ct), p(xi), p(w))
        # the hex27 / pyramid constants are truncated in the reference
        assert abs(w.sum() - vol) < (1e-11 if ct in (O.HEX27, O.HEX20, O.PYRAMID5) else 1e-14)
        for q in range(g):
            N, dN = np.zeros(n), np.zeros(3 * n)
            L.oracle_shape(C.c_int(ct), p(np.ascontiguousarray(xi[3 * q:3 * q + 3])), p(N), p(dN))
            assert abs(N.sum() - 1) < 1e-14 and np.abs(dN.reshape(n, 3).sum(0)).max() < 1e-13
        # Kronecker property at the nodes (the reference's "node" points of wedge6 are the first six of the wedge18 table,
        # not the places where the wedge6 functions are 1: 4C_fem_general_utils_local_connectivity_matrices.hpp:564-575)
        if ct == O.WEDGE6:
            continue
        xn = np.zeros(3 * n)
        L.oracle_node_param_coords(C.c_int(ct), p(xn))
        for a in range(n):
            N = np.zeros(n)
            L.oracle_shape(C.c_int(ct), p(np.ascontiguousarray(xn[3 * a:3 * a + 3])), p(N), None)
            e = np.zeros(n)
            e[a] = 1
            np.testing.assert_allclose(N, e, atol=1e-14)
