"""fc_solid_evaluate_stress_noalloc (include/fourc_b200.h): the allocation-free nodal-stress entry point against the
existing fc_solid_evaluate_stress, through the C ABI.  The entry point has no counterpart in the reference;
its scratch is one fixed device array inside the library that keeps the previous
call's contents, so the checks run on several calls in a row with different cell types and states.

Collected last (file name) so that a failure here cannot hide another test under `pytest -x`."""
import gc

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
TOL = 1e-12  # relative to max|stress|, the tolerance of the other stress tests


def _u(X, amp, phase):
    """smooth displacement with det F well inside (0.7, 1.4)"""
    u = np.zeros_like(X)
    u[:, 0] = amp * np.sin(2 * np.pi * X[:, 1] + phase) + 0.02 * X[:, 0]
    u[:, 1] = amp * np.cos(2 * np.pi * X[:, 2] - phase) - 0.01 * X[:, 1]
    u[:, 2] = amp * np.sin(2 * np.pi * X[:, 0] * X[:, 1] + phase) + 0.015 * X[:, 2]
    return np.ascontiguousarray(u.reshape(-1))


def _context(fc, ct, n, form, mat, young, nue):
    g = fc.Grid(ct, n)
    s = fc.Solid(0)
    s.set_mesh(ct, form, g.X, g.node_gid, g.node_owner, g.conn, g.ele_gid)
    s.set_material(mat, young, nue)
    return g, s


def _compare(fc, torch, g, s, u, stress_type):
    _, ref = s.evaluate_stress(u, stress_type, want_gp=False)            # existing entry point (allocates its scratch)
    u_dev = torch.from_numpy(u).cuda()
    out = torch.full((g.n_row_nodes, 6), float("nan"), dtype=torch.float64, device="cuda")
    s.evaluate_stress_noalloc(u_dev, out, stress_type)                   # new entry point, asynchronous
    torch.cuda.synchronize()
    first = out.cpu().numpy()
    # the same call again with the free device memory watched: it must not move.  (Watched on a repeated call so that
    # CUDA's lazy loading of a kernel at its first launch is not mistaken for an allocation; contexts of earlier tests that
    # still wait for the garbage collector would free theirs inside the window, so collect first and keep the collector out.)
    out.fill_(float("nan"))
    gc.collect()
    torch.cuda.synchronize()
    gc.disable()
    try:
        free0, _ = torch.cuda.mem_get_info()
        s.evaluate_stress_noalloc(u_dev, out, stress_type)
        torch.cuda.synchronize()
        free1, _ = torch.cuda.mem_get_info()
    finally:
        gc.enable()
    assert free1 == free0, "the no-alloc entry point changed the device's free memory"
    assert np.all(np.isfinite(first)) and np.max(np.abs(first - ref)) <= TOL * np.max(np.abs(ref))
    got = out.cpu().numpy()
    scale = np.max(np.abs(ref))
    assert scale > 0 and np.all(np.isfinite(got))
    err = np.max(np.abs(got - ref)) / scale
    assert err <= TOL, err
    return err


def test_noalloc_matches_existing_entry_point_on_calls_in_a_row(fc):
    torch = pytest.importorskip("torch")
    cases = [
        (fc.FC_HEX27, (4, 3, 3), fc.FC_FORM_DISPLACEMENT, fc.FC_MAT_ELASTHYPER_COUPNEOHOOKE, 10.0, 0.25),  # fills 27x27 of the matrix slot
        (fc.FC_HEX8, (7, 6, 5), fc.FC_FORM_DISPLACEMENT, fc.FC_MAT_ELASTHYPER_COUPNEOHOOKE, 10.0, 0.25),   # smaller: stale contents behind it
        (fc.FC_HEX8, (5, 5, 4), fc.FC_FORM_FBAR, fc.FC_MAT_ELASTHYPER_COUPNEOHOOKE, 10.0, 0.45),
        (fc.FC_HEX8, (6, 4, 4), fc.FC_FORM_DISPLACEMENT, fc.FC_MAT_STVENANTKIRCHHOFF, 210.0, 0.3),
    ]
    for ct, n, form, mat, young, nue in cases:
        g, s = _context(fc, ct, n, form, mat, young, nue)
        # two calls in a row on the same context with different states and stress types, then the first state again
        _compare(fc, torch, g, s, _u(g.X, 0.010, 0.3), fc.FC_STRESS_CAUCHY)
        _compare(fc, torch, g, s, _u(g.X, 0.015, 1.1), fc.FC_STRESS_2PK)
        _compare(fc, torch, g, s, _u(g.X, 0.010, 0.3), fc.FC_STRESS_CAUCHY)
        s.close()


def test_noalloc_rejects_sizes_beyond_the_fixed_scratch(fc):
    torch = pytest.importorskip("torch")
    g, s = _context(fc, fc.FC_HEX8, (32, 32, 32), fc.FC_FORM_DISPLACEMENT, fc.FC_MAT_ELASTHYPER_COUPNEOHOOKE, 10.0, 0.25)
    u_dev = torch.zeros(3 * g.n_col_nodes, dtype=torch.float64, device="cuda")
    out = torch.zeros((g.n_row_nodes, 6), dtype=torch.float64, device="cuda")
    with pytest.raises(fc.capi.FourCError) as e:
        s.evaluate_stress_noalloc(u_dev, out)
    assert e.value.code == fc.FC_ERR_INVALID and "FC_NOALLOC_MAX" in str(e.value)
    # host buffers are refused as well (staging them would allocate)
    g2, s2 = _context(fc, fc.FC_HEX8, (3, 3, 3), fc.FC_FORM_DISPLACEMENT, fc.FC_MAT_ELASTHYPER_COUPNEOHOOKE, 10.0, 0.25)
    with pytest.raises(fc.capi.FourCError):
        s2.evaluate_stress_noalloc(np.zeros(3 * g2.n_col_nodes), np.zeros((g2.n_row_nodes, 6)))
    s.close()
    s2.close()
