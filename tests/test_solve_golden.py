"""Solve-level golden vectors of the reference (RESULT DESCRIPTION sections), reproduced
with the mock Newton driver on top of (a) the CPU oracle and (b) the CUDA library."""
import numpy as np
import pytest
from newton_driver import face_traction_forces, hex8_faces_on_plane, newton_statics, solve_with_dirichlet

HEX8, DISP, SVK, NH = 0, 0, 0, 1


def altgeo_problem(oracle):
    """tests/input_files/sohex8_disp_altgeogeneration.dat: STRUCTURE DOMAIN [0,4]^3, INTERVALS 3 3 3,
    HEX8, MAT_Struct_StVenantKirchhoff YOUNG 1 NUE 0, x- face clamped, x+ face Neumann 1.0 in x,
    statics, MAXITER 1.  RESULT: NODE 49 (gid 48) dispx 4.0 dispy 0 dispz 0, TOLERANCE 1e-14."""
    job = oracle.Job.cuboid(HEX8, (3, 3, 3), lo=(0, 0, 0), hi=(4, 4, 4))
    job.set_problem(DISP, SVK, 1.0, 0.0)
    idx = {int(g): i for i, g in enumerate(job.node_gid)}
    conn = np.vectorize(idx.get)(job.conn)
    fixed_nodes = np.where(np.abs(job.X[:, 0]) < 1e-12)[0]
    fixed = (3 * fixed_nodes[:, None] + np.arange(3)).reshape(-1)
    faces = hex8_faces_on_plane(job.X, conn, 0, 4.0)
    f_ext = face_traction_forces(job.X, faces, np.array([1.0, 0.0, 0.0])).reshape(-1)
    return job, conn, fixed, f_ext, idx[48]


def test_altgeogeneration_oracle(oracle):
    job, conn, fixed, f_ext, n48 = altgeo_problem(oracle)
    o = job.rank(0)
    assert np.array_equal(o["col_node_gid"], job.node_gid)  # 1 rank: row layout == mesh order
    job.set_u_nodal(np.zeros_like(job.X))
    job.evaluate(oracle.ACTION_NLNSTIFF)
    K = job.get_K(0, o["nnz"])
    du = solve_with_dirichlet(o["rowptr"], o["colind"], K, f_ext, fixed)
    np.testing.assert_allclose(du[3 * n48:3 * n48 + 3], [4.0, 0.0, 0.0], atol=1e-12, rtol=0)


def test_altgeogeneration_np2_partition(oracle):
    """the reference runs this input with NP 2: the structured split must leave the result unchanged and
    node 49 must exist with the generator's lattice numbering"""
    job = oracle.Job.cuboid(HEX8, (3, 3, 3), lo=(0, 0, 0), hi=(4, 4, 4), nranks=2)
    assert job.subdivisions() == (2, 1, 1)
    assert 48 in set(int(g) for g in job.node_gid)
    i = list(job.node_gid).index(48)
    np.testing.assert_allclose(job.X[i], [4.0, 4.0, 0.0])
    assert sorted(set(job.ele_owner)) == [0, 1]


def test_newton_converges_quadratically_oracle(oracle):
    """hex8 Neo-Hooke (YOUNG 10 NUE 0.25, the reference's test material) cube under traction: the
    residual norms of a full Newton solve must fall quadratically iff K = d f_int / d u."""
    job = oracle.Job.cuboid(HEX8, (4, 4, 4))
    job.set_problem(DISP, NH, 10.0, 0.25)
    o = job.rank(0)
    idx = {int(g): i for i, g in enumerate(job.node_gid)}
    conn = np.vectorize(idx.get)(job.conn)
    fixed_nodes = np.where(np.abs(job.X[:, 0]) < 1e-12)[0]
    fixed = (3 * fixed_nodes[:, None] + np.arange(3)).reshape(-1)
    f_ext = face_traction_forces(job.X, hex8_faces_on_plane(job.X, conn, 0, 1.0), np.array([1.0, 0.3, 0.0])).reshape(-1)

    def assemble(u):
        job.set_u_nodal(u.reshape(-1, 3))
        bad, _ = job.evaluate(oracle.ACTION_NLNSTIFF)
        assert bad == 0
        return job.get_f(0, o["n_row_dof"]), job.get_K(0, o["nnz"])

    u, norms = newton_statics(assemble, o["n_row_dof"], f_ext, fixed, o["rowptr"], o["colind"])
    assert norms[-1] < 1e-10 and len(norms) <= 8
    assert norms[-2] < 1e-4 and norms[-1] < 10 * norms[-2] ** 2 + 1e-13  # quadratic tail
    np.save("/tmp/_newton_norms_oracle.npy", np.array(norms))


@pytest.mark.gpu
def test_altgeogeneration_cuda(fc, oracle):
    job, conn, fixed, f_ext, n48 = altgeo_problem(oracle)
    grid = fc.Grid(fc.FC_HEX8, (3, 3, 3), lo=(0, 0, 0), hi=(4, 4, 4))
    assert np.array_equal(grid.node_gid, job.node_gid) and np.array_equal(grid.X, job.X)
    s = fc.Solid(0)
    s.set_mesh(fc.FC_HEX8, fc.FC_FORM_DISPLACEMENT, grid.X, grid.node_gid, grid.node_owner, grid.conn, grid.ele_gid)
    s.set_material(fc.FC_MAT_STVENANTKIRCHHOFF, 1.0, 0.0)
    n_row, n_col, nnz = s.build_graph()
    rowptr, colind = s.get_graph()
    K, f = np.zeros(nnz), np.zeros(n_row)
    assert s.evaluate(fc.FC_ACTION_NLNSTIFF, np.zeros(n_col), f, K, flags=fc.FC_EVAL_ZERO_FIRST) == (0, -1)
    du = solve_with_dirichlet(rowptr, colind, K, f_ext, fixed)
    np.testing.assert_allclose(du[3 * n48:3 * n48 + 3], [4.0, 0.0, 0.0], atol=1e-12, rtol=0)
    s.close()


@pytest.mark.gpu
@pytest.mark.parametrize("flags", [0, 2, 4])
def test_newton_iteration_counts_and_norms_match_oracle(fc, oracle, flags):
    """north_star: 'Newton iteration counts and residual norms of the full solve must match'."""
    n = (4, 4, 4)
    job = oracle.Job.cuboid(HEX8, n)
    job.set_problem(DISP, NH, 10.0, 0.25)
    o = job.rank(0)
    idx = {int(g): i for i, g in enumerate(job.node_gid)}
    conn = np.vectorize(idx.get)(job.conn)
    fixed_nodes = np.where(np.abs(job.X[:, 0]) < 1e-12)[0]
    fixed = (3 * fixed_nodes[:, None] + np.arange(3)).reshape(-1)
    f_ext = face_traction_forces(job.X, hex8_faces_on_plane(job.X, conn, 0, 1.0), np.array([1.0, 0.3, 0.0])).reshape(-1)

    def assemble_oracle(u):
        job.set_u_nodal(u.reshape(-1, 3))
        job.evaluate(oracle.ACTION_NLNSTIFF)
        return job.get_f(0, o["n_row_dof"]), job.get_K(0, o["nnz"])

    grid = fc.Grid(fc.FC_HEX8, n)
    s = fc.Solid(0)
    s.set_mesh(fc.FC_HEX8, fc.FC_FORM_DISPLACEMENT, grid.X, grid.node_gid, grid.node_owner, grid.conn, grid.ele_gid)
    s.set_material(fc.FC_MAT_ELASTHYPER_COUPNEOHOOKE, 10.0, 0.25)
    n_row, n_col, nnz = s.build_graph()
    s.set_dirichlet_rows(fixed)  # rows skipped by the scatter like dbcmaps_ (4C_linalg_sparsematrix.cpp:480)

    def assemble_cuda(u):
        K, f = np.zeros(nnz), np.zeros(n_row)
        assert s.evaluate(fc.FC_ACTION_NLNSTIFF, u, f, K, flags=flags | fc.FC_EVAL_ZERO_FIRST) == (0, -1)
        # Dirichlet rows were skipped: they must still be exactly zero
        rp = o["rowptr"]
        for d in fixed[:6]:
            assert not K[rp[d]:rp[d + 1]].any()
        return f, K

    u0, n0 = newton_statics(assemble_oracle, n_row, f_ext, fixed, o["rowptr"], o["colind"])
    u1, n1 = newton_statics(assemble_cuda, n_row, f_ext, fixed, o["rowptr"], o["colind"])
    assert len(n0) == len(n1)
    for a, b in zip(n0, n1):
        assert abs(a - b) <= 1e-10 * max(n0[0], 1e-300) + 1e-13
    assert np.max(np.abs(u0 - u1)) <= 1e-11 * np.max(np.abs(u0))
    s.close()
