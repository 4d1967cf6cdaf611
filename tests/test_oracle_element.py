"""Element-level K_e / f_e of the oracle are pinned by no reference unit test: they are
cross-checked here by properties the reference's formulation must satisfy."""
import numpy as np
import oracle_lib as O
import pytest

CASES = [(O.HEX8, O.FORM_DISP, O.MAT_NEOHOOKE, 0.25), (O.HEX8, O.FORM_DISP, O.MAT_SVK, 0.3),
         (O.HEX8, O.FORM_FBAR, O.MAT_NEOHOOKE, 0.45), (O.HEX27, O.FORM_DISP, O.MAT_NEOHOOKE, 0.25),
         (O.TET10, O.FORM_DISP, O.MAT_NEOHOOKE, 0.25), (O.TET10, O.FORM_DISP, O.MAT_SVK, 0.3)]


def element_nodes(ct, rng, distort=0.05):
    import ctypes as C
    n = O.NEN[ct]
    xi = np.zeros(3 * n)
    O.lib.oracle_node_param_coords(C.c_int(ct), xi.ctypes.data_as(C.c_void_p))
    X = xi.reshape(n, 3) * (0.5 if ct != O.TET10 else 1.0)
    return X + distort * rng.uniform(-1, 1, X.shape) * 0.5


@pytest.mark.parametrize("ct,form,mat,nue", CASES)
def test_tangent_is_derivative_of_force(ct, form, mat, nue):
    rng = np.random.default_rng(7)
    X = element_nodes(ct, rng)
    u = 0.03 * rng.standard_normal(X.shape)
    err, fe, Ke, _ = O.element_evaluate(ct, form, mat, 10.0, nue, X, u)
    assert err == 0
    nd, h = fe.size, 1e-6
    Kfd = np.zeros((nd, nd))
    for k in range(nd):
        du = np.zeros(nd)
        du[k] = h
        _, fp, _, _ = O.element_evaluate(ct, form, mat, 10.0, nue, X, u + du.reshape(-1, 3), want_K=False)
        _, fm, _, _ = O.element_evaluate(ct, form, mat, 10.0, nue, X, u - du.reshape(-1, 3), want_K=False)
        Kfd[:, k] = (fp - fm) / (2 * h)
    assert np.max(np.abs(Ke - Kfd)) <= 2e-7 * np.max(np.abs(Ke))
    if form == O.FORM_DISP:
        assert np.max(np.abs(Ke - Ke.T)) <= 1e-13 * np.max(np.abs(Ke))  # hyperelastic + conservative


@pytest.mark.parametrize("ct,form,mat,nue", CASES)
def test_rigid_body_motion_is_force_free(ct, form, mat, nue):
    rng = np.random.default_rng(3)
    X = element_nodes(ct, rng)
    a = 0.4
    R = np.array([[np.cos(a), -np.sin(a), 0], [np.sin(a), np.cos(a), 0], [0, 0, 1.0]])
    u = X @ R.T - X + np.array([0.3, -0.2, 0.1])
    err, fe, Ke, _ = O.element_evaluate(ct, form, mat, 10.0, nue, X, u)
    assert err == 0 and np.max(np.abs(fe)) < 1e-13
    # translations are in the null space of K
    for d in range(3):
        t = np.zeros((O.NEN[ct], 3))
        t[:, d] = 1.0
        assert np.max(np.abs(Ke @ t.reshape(-1))) <= 1e-12 * np.max(np.abs(Ke))


def test_negative_jacobian_is_reported():
    rng = np.random.default_rng(1)
    X = element_nodes(O.HEX8, rng, 0.0)
    X[[0, 1]] = X[[1, 0]]  # tangled element
    err, _, _, _ = O.element_evaluate(O.HEX8, O.FORM_DISP, O.MAT_SVK, 1.0, 0.3, X, np.zeros_like(X))
    assert err == 1


def test_mass_matrix_integrates_density():
    rng = np.random.default_rng(5)
    X = element_nodes(O.HEX8, rng, 0.0)
    _, _, _, Me = O.element_evaluate(O.HEX8, O.FORM_DISP, O.MAT_SVK, 1.0, 0.3, X, np.zeros_like(X), want_M=True, rho=2.5)
    assert abs(Me[0::3, 0::3].sum() - 2.5 * 1.0) < 1e-13  # unit cube
