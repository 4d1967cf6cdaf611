"""The oracle's NURBS basis (oracle/fc_oracle_nurbs.hpp) and its nurbs27 element: properties that pin the restatement without
a GPU - partition of unity, derivatives against differences, Bernstein polynomials for an open knot vector with unit weights,
exact reproduction of an affine field, K = df/du.  The solve-level pins are the reference's own results
(solid_ele_nurbs27_Standard_Static, solid_runtime_nurbs in test_golden_reference_solves.py)."""
import ctypes as C

import numpy as np
import oracle_lib as o

NURBS27 = o.NURBS27
KN = np.array([[0, 0, 0, 1, 1, 1], [0, 0, 0.3, 0.6, 1, 1], [0, 0.2, 0.5, 0.75, 1, 1.0]])


def basis(xi):
    N, dN = o.shape(NURBS27, xi), np.zeros(81)
    o.lib.oracle_shape(C.c_int(NURBS27), o._p(o._d(xi)), None, o._p(dN))
    return N, dN.reshape(27, 3)


def test_partition_of_unity_and_derivatives():
    rng = np.random.default_rng(0)
    o.set_nurbs_element(KN, rng.uniform(0.5, 1.5, 27))
    for xi in rng.uniform(-1, 1, (5, 3)):
        N, dN = basis(xi)
        assert abs(N.sum() - 1.0) < 1e-14 and N.min() >= 0.0 and np.max(np.abs(dN.sum(0))) < 1e-13
        h = 1e-6
        for d in range(3):
            e = np.zeros(3)
            e[d] = h
            fd = (o.shape(NURBS27, xi + e) - o.shape(NURBS27, xi - e)) / (2 * h)
            assert np.max(np.abs(fd - dN[:, d])) < 1e-9


def test_open_knot_vector_with_unit_weights_is_bernstein():
    o.set_nurbs_element(np.tile([0, 0, 0, 1, 1, 1.0], (3, 1)), np.ones(27))
    xi = np.array([0.3, -0.2, 0.7])
    t = (xi + 1) / 2
    B = [np.array([(1 - s) ** 2, 2 * s * (1 - s), s ** 2]) for s in t]
    ref = np.array([B[0][a % 3] * B[1][(a // 3) % 3] * B[2][a // 9] for a in range(27)])
    assert np.max(np.abs(o.shape(NURBS27, xi) - ref)) < 1e-15


def test_affine_field_and_tangent():
    rng = np.random.default_rng(1)
    w = rng.uniform(0.5, 1.5, 27)
    o.set_nurbs_element(KN, w)
    # control points at the Greville abscissae of the three knot vectors: the geometry map is the identity on the knot span
    g = [np.array([(k[i + 1] + k[i + 2]) / 2 for i in range(3)]) for k in KN]
    X = np.array([[g[0][a % 3], g[1][(a // 3) % 3], g[2][a // 9]] for a in range(27)])
    # rational basis: an affine displacement of the control points is an affine displacement field -> constant stress
    H = 0.05 * rng.uniform(-1, 1, (3, 3))
    u = X @ H.T
    err, S = o.element_gp_stress(NURBS27, 0, 0, 10.0, 0.3, X, u, 0)
    assert err == 0 and np.max(np.abs(S - S[0])) < 1e-13 * np.max(np.abs(S))
    F = np.eye(3) + H
    E = 0.5 * (F.T @ F - np.eye(3))
    lam, mu = 10.0 * 0.3 / (1.3 * 0.4), 10.0 / 2.6
    Sref = lam * np.trace(E) * np.eye(3) + 2 * mu * E
    assert np.max(np.abs(S[0] - [Sref[0, 0], Sref[1, 1], Sref[2, 2], Sref[0, 1], Sref[1, 2], Sref[0, 2]])) < 1e-13
    # K = df/du on a distorted element
    Xd = X + rng.uniform(-0.02, 0.02, X.shape)
    ud = rng.uniform(-0.02, 0.02, X.shape)
    err, f, K, _ = o.element_evaluate(NURBS27, 0, 1, 10.0, 0.3, Xd, ud)
    assert err == 0 and np.max(np.abs(K - K.T)) < 1e-14 * np.max(np.abs(K))
    h, Kfd = 1e-6, np.zeros((81, 81))
    for j in range(81):
        e = np.zeros(81)
        e[j] = h
        fp = o.element_evaluate(NURBS27, 0, 1, 10.0, 0.3, Xd, ud + e.reshape(27, 3), want_K=False)[1]
        fm = o.element_evaluate(NURBS27, 0, 1, 10.0, 0.3, Xd, ud - e.reshape(27, 3), want_K=False)[1]
        Kfd[:, j] = (fp - fm) / (2 * h)
    assert np.max(np.abs(K - Kfd)) < 1e-8 * np.max(np.abs(K))


def test_face_basis():
    rng = np.random.default_rng(2)
    o.set_nurbs_element(KN[[0, 2]], rng.uniform(0.5, 1.5, 9))
    x = np.array([[i, j, 0.1 * i * j] for j in range(3) for i in range(3)], float)
    fe, gc = o.surface_neumann(o.NURBS9, x[None], [1, 1, 1], [1.0, 2.0, -1.0], want_coords=True)
    area = fe[0].sum(0) / [1.0, 2.0, -1.0]
    assert np.max(np.abs(area - area[0])) < 1e-13 and area[0] > 0 and gc.shape == (1, 9, 3)
