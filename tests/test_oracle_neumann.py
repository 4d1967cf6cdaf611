"""Properties of the oracle's orthopressure load (oracle/fc_oracle_neumann.hpp) that hold without a reference value."""
import numpy as np


def test_orthopressure_on_a_closed_surface_has_no_resultant():
    """uniform pressure on the six faces of a distorted hex8 (normals from the outward-oriented face numbering): the forces sum
    to zero (divergence theorem), and on a single flat face to pressure x area along its normal"""
    import oracle_lib as o
    rng = np.random.default_rng(0)
    xi = o.node_param_coords(o.HEX8)
    X = xi + 0.2 * (rng.random(xi.shape) - 0.5)
    faces = np.array([[0, 3, 2, 1], [0, 1, 5, 4], [1, 2, 6, 5], [2, 3, 7, 6], [0, 4, 7, 3], [4, 5, 6, 7]])
    fe = o.surface_neumann_orthopressure(o.QUAD4, X[faces], 2.5)
    assert np.max(np.abs(fe.sum(axis=(0, 1)))) < 1e-14
    flat = np.array([[0.0, 0, 1], [2, 0, 1], [2, 3, 1], [0, 3, 1]])
    f1 = o.surface_neumann_orthopressure(o.QUAD4, flat[None], -0.5)
    assert np.allclose(f1.sum(axis=(0, 1)), [0.0, 0.0, -0.5 * 6.0], atol=1e-14)


def test_orthopressure_linearization_is_minus_the_derivative_of_the_force():
    import oracle_lib as o
    rng = np.random.default_rng(1)
    for ft, nn in ((o.QUAD4, 4), (o.QUAD8, 8), (o.QUAD9, 9), (o.TRI6, 6), (o.TRI3, 3), (o.QUAD6, 6)):
        x = rng.random((1, nn, 3))
        fac = 1.0 + rng.random((1, o.FACE_NGP[ft], 3))
        K = o.surface_neumann_orthopressure_linearization(ft, x, 1.3, fac)[0]
        h, fd = 1e-6, np.zeros((3 * nn, 3 * nn))
        for j in range(3 * nn):
            e = np.zeros(3 * nn)
            e[j] = h
            fp = o.surface_neumann_orthopressure(ft, x + e.reshape(1, nn, 3), 1.3, fac)
            fm = o.surface_neumann_orthopressure(ft, x - e.reshape(1, nn, 3), 1.3, fac)
            fd[:, j] = ((fp - fm) / (2 * h)).reshape(-1)
        assert np.max(np.abs(K + fd)) < 1e-8 * max(np.max(np.abs(K)), 1e-3), ft
