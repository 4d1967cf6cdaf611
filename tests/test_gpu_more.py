"""More GPU tests through the C ABI: device-pointer path, multi-rank decomposition emulated on one GPU
(one context per rank, halo pack kernel + peer copies), error reporting, full-size invariants."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
HEX8, DISP, FBAR, SVK, NH = 0, 0, 1, 0, 1


def field(X, h):
    A = np.array([[0.03, 0.01, -0.02], [0.00, -0.02, 0.015], [0.01, 0.02, 0.025]])
    return X @ A.T + 0.01 * np.sin(2 * np.pi * X[:, [1, 2, 0]])


def make_ctx(fc, grid, form=DISP, mat=NH, nue=0.25, rank=0, nranks=1):
    s = fc.Solid(0, rank, nranks)
    s.set_mesh(fc.FC_HEX8, form, grid.X, grid.node_gid, grid.node_owner, grid.conn, grid.ele_gid)
    s.set_material(mat, 10.0, nue)
    s.build_graph()
    return s


@pytest.mark.parametrize("nranks,n", [(2, (8, 6, 5)), (4, (8, 8, 4)), (8, (6, 6, 6))])
def test_multi_rank_on_one_gpu_matches_oracle(fc, oracle, nranks, n):
    """Variant A of SURVEY 8e: halo import of u, redundant ghost elements, owner rows only."""
    import torch
    job = oracle.Job.cuboid(HEX8, n, nranks=nranks)
    job.set_problem(DISP, NH, 10.0, 0.25)
    u_nodes = field(job.X, 1.0 / max(n))
    job.set_u_nodal(u_nodes)
    assert job.evaluate(oracle.ACTION_NLNSTIFF, max_threads=4)[0] == 0
    grids = [fc.Grid(HEX8, n, nranks=nranks, rank=r) for r in range(nranks)]
    ctxs = [make_ctx(fc, grids[r], rank=r, nranks=nranks) for r in range(nranks)]
    dev = torch.device("cuda", 0)
    # row-layout state per rank on the device; ghost part poisoned
    u_col, sendbuf = [], []
    for r, g in enumerate(grids):
        u = torch.full((3 * g.n_col_nodes,), float("nan"), dtype=torch.float64, device=dev)
        u[:3 * g.n_row_nodes] = torch.from_numpy(field(g.X[:g.n_row_nodes], 0).reshape(-1)).to(dev)
        u_col.append(u)
        lids = (3 * g.send_nodes[:, None] + np.arange(3)).reshape(-1)
        ctxs[r].halo_set_send(lids)
        sendbuf.append(torch.empty(len(lids), dtype=torch.float64, device=dev))
        ctxs[r].halo_pack(u, sendbuf[r])
    torch.cuda.synchronize()
    # the exchange (NCCL all-to-all in bench.py) as device copies: p's message for q lands behind q's owned DOFs
    for q, gq in enumerate(grids):
        off = 3 * gq.n_row_nodes
        for p in range(nranks):
            cnt = 3 * int(gq.recv_count[p])
            s0 = 3 * int(grids[p].send_count[:q].sum())
            u_col[q][off:off + cnt] = sendbuf[p][s0:s0 + cnt]
            off += cnt
    for r, g in enumerate(grids):
        assert not torch.isnan(u_col[r]).any()
        n_row, n_col, nnz = ctxs[r].graph_sizes()
        K = torch.zeros(nnz, dtype=torch.float64, device=dev)
        f = torch.zeros(n_row, dtype=torch.float64, device=dev)
        assert ctxs[r].evaluate(fc.FC_ACTION_NLNSTIFF, u_col[r], f, K, flags=fc.FC_EVAL_ZERO_FIRST) == (0, -1)
        o = job.rank(r)
        Kr, fr = job.get_K(r, o["nnz"]), job.get_f(r, o["n_row_dof"])
        if nnz:
            assert np.max(np.abs(K.cpu().numpy() - Kr)) <= 1e-12 * np.max(np.abs(Kr))
            assert np.max(np.abs(f.cpu().numpy() - fr)) <= 1e-12 * np.max(np.abs(Kr)) / max(n)
        ctxs[r].close()


def test_negative_jacobian_and_bad_state_are_reported(fc):
    g = fc.Grid(HEX8, (4, 3, 3))
    conn = g.conn.copy()
    conn[5, [0, 1]] = conn[5, [1, 0]]  # tangle element 5
    s = fc.Solid(0)
    s.set_mesh(fc.FC_HEX8, DISP, g.X, g.node_gid, g.node_owner, conn, g.ele_gid)
    s.set_material(NH, 10.0, 0.25)
    n_row, n_col, nnz = s.build_graph()
    K, f = np.zeros(nnz), np.zeros(n_row)
    cnt, gid = s.evaluate(fc.FC_ACTION_NLNSTIFF, np.zeros(n_col), f, K)
    assert cnt == 1 and gid == int(g.ele_gid[5])
    s.close()
    # inverted state (det F < 0 everywhere): disp-based Neo-Hooke sees I3 > 0 and must NOT flag it (as the reference),
    # a collapsed state (F = 0) must
    s = make_ctx(fc, g)
    n_row, n_col, nnz = s.graph_sizes()
    u = (-g.X).reshape(-1).copy()  # x = 0 for all nodes
    cnt, gid = s.evaluate(fc.FC_ACTION_NLNSTIFF, u, np.zeros(n_row), np.zeros(nnz))
    assert cnt == g.n_col_eles and gid == int(g.ele_gid.min())
    s.close()


def test_argument_errors(fc):
    g = fc.Grid(HEX8, (2, 2, 2))
    s = fc.Solid(0)
    with pytest.raises(fc.FourCError):
        s.evaluate(fc.FC_ACTION_NLNSTIFF, np.zeros(3), np.zeros(3), np.zeros(3))  # no mesh
    with pytest.raises(fc.FourCError):
        s.set_mesh(fc.FC_HEX27, FBAR, g.X, g.node_gid, g.node_owner, g.conn, g.ele_gid)  # F-bar is hex8 only
    s.set_mesh(fc.FC_HEX8, DISP, g.X, g.node_gid, g.node_owner, g.conn, g.ele_gid)
    with pytest.raises(fc.FourCError):
        s.set_material(NH, 10.0, 0.5)
    s.set_material(NH, 10.0, 0.25)
    with pytest.raises(fc.FourCError):
        s.evaluate(fc.FC_ACTION_NLNSTIFF, np.zeros(81), np.zeros(81), np.zeros(10))  # no graph yet
    # a graph that misses couplings is rejected
    rowptr = np.arange(0, 82 * 3, 3, dtype=np.int64)[:82]
    colind = np.repeat(np.arange(27) * 3, 3 * 3).reshape(27, 3, 3) + np.arange(3)
    with pytest.raises(fc.FourCError) as ei:
        s.set_graph(rowptr, colind.reshape(-1))
    assert ei.value.code == fc.FC_ERR_GRAPH
    s.close()


def test_c2_full_size_invariants(fc):
    """BASELINE C2 (100^3): size-independent properties of the assembled system on the device."""
    import torch
    n = (100, 100, 100)
    g = fc.Grid(HEX8, n)
    s = make_ctx(fc, g)
    n_row, n_col, nnz = s.graph_sizes()
    assert nnz == 9 * 301 ** 3 == 245_438_109
    dev = torch.device("cuda", 0)
    u = torch.from_numpy(field(g.X, 0.01).reshape(-1)).to(dev)
    K = torch.empty(nnz, dtype=torch.float64, device=dev)
    f = torch.empty(n_row, dtype=torch.float64, device=dev)
    assert s.evaluate(fc.FC_ACTION_NLNSTIFF, u, f, K, flags=fc.FC_EVAL_ZERO_FIRST) == (0, -1)
    rowptr, colind = s.get_graph()
    rp = torch.from_numpy(rowptr).to(dev)
    ci = torch.from_numpy(colind).to(dev).long()
    # (1) translations are in the null space of K: row sums over each displacement component vanish
    rows = torch.repeat_interleave(torch.arange(n_row, device=dev), (rp[1:] - rp[:-1]))
    kmax = K.abs().max()
    for d in range(3):
        m = (ci % 3) == d
        rs = torch.zeros(n_row, dtype=torch.float64, device=dev).index_add_(0, rows[m], K[m])
        assert rs.abs().max() <= 1e-12 * kmax
    # (2) symmetry: sum_ij K_ij x_i y_j == sum_ij K_ij y_i x_j for random x, y
    gen = torch.Generator(device=dev).manual_seed(1)
    x = torch.rand(n_row, dtype=torch.float64, device=dev, generator=gen)
    y = torch.rand(n_row, dtype=torch.float64, device=dev, generator=gen)
    a = (K * x[rows] * y[ci]).sum()
    b = (K * y[rows] * x[ci]).sum()
    assert abs(a - b) <= 1e-11 * (K.abs() * x[rows] * y[ci]).sum()
    # (3) total internal force vanishes (self-equilibrated) and K dx ~ f(u+dx) - f(u)
    assert f.view(-1, 3).sum(0).abs().max() <= 1e-10 * f.abs().max() * 1e3
    dx = 1e-7 * (torch.rand(n_col, dtype=torch.float64, device=dev, generator=gen) - 0.5)
    f2 = torch.empty_like(f)
    s.evaluate(fc.FC_ACTION_INTERNALFORCE, u + dx, f2, None, flags=fc.FC_EVAL_ZERO_FIRST)
    Kdx = torch.zeros(n_row, dtype=torch.float64, device=dev).index_add_(0, rows, K * dx[ci])
    assert (f2 - f - Kdx).abs().max() <= 1e-5 * Kdx.abs().max()
    # (4) the owner-computed kernel, the FMA closed form and the generic kernel agree with the default at full size
    K2 = torch.empty_like(K)
    for flags in (fc.FC_EVAL_OWNER_COMPUTED, fc.FC_EVAL_CLOSED_FORM_FMA, fc.FC_EVAL_FORCE_GENERIC):
        s.evaluate(fc.FC_ACTION_NLNSTIFF, u, f2, K2, flags=flags | fc.FC_EVAL_ZERO_FIRST)
        assert (K2 - K).abs().max() <= 1e-12 * kmax
        assert (f2 - f).abs().max() <= 1e-12 * kmax * 0.01
    s.close()


@pytest.mark.parametrize("ct,n", [(0, (4, 3, 2)), (1, (2, 2, 2)), (2, None)])
def test_mass_matrix_parity(fc, oracle, ct, n):
    """consistent mass (struct_calc_nlnstiffmass / add_mass_matrix, tet10 with the 11-point rule) against the
    oracle's element mass matrices assembled on the host; rows sum to the nodal share of the total mass."""
    import structure_driver as sd
    if ct == 2:
        p = sd.Problem(sd.load("solid_ele_tet10_Standard_dynamic"), oracle.node_param_coords)
        X, gid, conn, egid = p.X, p.node_gid, p.conn, p.ele_gid
    else:
        g = fc.Grid(ct, n)
        X, gid, conn, egid = g.X * np.array([1.0, 1.3, 0.7]), g.node_gid, g.conn, g.ele_gid
    rho = 0.37
    s = fc.Solid(0)
    s.set_mesh(ct, DISP, X, gid, np.zeros(len(gid), np.int32), conn, egid)
    s.set_material(NH, 10.0, 0.25)
    n_row, n_col, nnz = s.build_graph()
    rowptr, colind = s.get_graph()
    M = np.full(nnz, 7.0)
    s.evaluate_mass(rho, M)
    Md = sd.csr_to_dense(rowptr, colind, M, n_row)
    Mo = np.zeros((n_row, n_row))
    for en in conn:
        err, _, _, Me = oracle.element_evaluate(ct, 0, 1, 10.0, 0.25, X[en], np.zeros((len(en), 3)), want_K=False, want_M=True, rho=rho)
        dofs = (3 * en[:, None] + np.arange(3)).reshape(-1)
        Mo[np.ix_(dofs, dofs)] += Me
    assert np.max(np.abs(Md - Mo)) <= 1e-13 * np.max(np.abs(Mo))
    # adding a second time doubles it (+= semantics without ZERO_FIRST)
    s.evaluate_mass(rho, M, flags=0)
    assert np.max(np.abs(sd.csr_to_dense(rowptr, colind, M, n_row) - 2 * Mo)) <= 1e-13 * np.max(np.abs(Mo))
    s.close()


@pytest.mark.parametrize("form,mat,nue", [(FBAR, NH, 0.499), (DISP, SVK, 0.3)])
def test_full_size_kernel_agreement_fbar_svk(fc, form, mat, nue):
    """BASELINE C4's element (F-bar Neo-Hooke nu=0.499) and C1's (SVK) at 10^6 elements: the tensor-core closed form
    and the generic B-operator kernel agree, translations are in the null space, f is self-equilibrated."""
    import torch
    n = (100, 100, 100)
    g = fc.Grid(HEX8, n)
    s = fc.Solid(0)
    s.set_mesh(fc.FC_HEX8, form, g.X, g.node_gid, g.node_owner, g.conn, g.ele_gid)
    s.set_material(mat, 10.0, nue)
    n_row, n_col, nnz = s.build_graph()
    dev = torch.device("cuda", 0)
    u = torch.from_numpy(field(g.X, 0.01).reshape(-1)).to(dev)
    K = torch.empty(nnz, dtype=torch.float64, device=dev)
    f = torch.empty(n_row, dtype=torch.float64, device=dev)
    assert s.evaluate(fc.FC_ACTION_NLNSTIFF, u, f, K, flags=fc.FC_EVAL_ZERO_FIRST) == (0, -1)
    kmax = K.abs().max()
    rowptr, colind = s.get_graph()
    rp = torch.from_numpy(rowptr).to(dev)
    ci = torch.from_numpy(colind).to(dev).long()
    rows = torch.repeat_interleave(torch.arange(n_row, device=dev), (rp[1:] - rp[:-1]))
    for d in range(3):
        m = (ci % 3) == d
        rs = torch.zeros(n_row, dtype=torch.float64, device=dev).index_add_(0, rows[m], K[m])
        assert rs.abs().max() <= 1e-12 * kmax
    assert f.view(-1, 3).sum(0).abs().max() <= 1e-10 * f.abs().max() * 1e3
    del rows, ci, rp
    K2, f2 = torch.empty_like(K), torch.empty_like(f)
    assert s.evaluate(fc.FC_ACTION_NLNSTIFF, u, f2, K2, flags=fc.FC_EVAL_FORCE_GENERIC | fc.FC_EVAL_ZERO_FIRST) == (0, -1)
    assert (K2 - K).abs().max() <= 1e-12 * kmax
    assert (f2 - f).abs().max() <= 1e-12 * kmax * 0.01
    s.close()


@pytest.mark.gpu
def test_element_order_and_kernel_choice_give_the_same_assembly(fc):
    """the caller's element order (here shuffled) changes the order of the reductions of the per-element kernel and the
    contents of the brick plan, not the result: rounding only.  The owner-computed kernel is bit-reproducible."""
    g = fc.Grid(HEX8, (15, 9, 8))
    u = field(g.X, 0).reshape(-1).copy()
    rng = np.random.default_rng(3)
    res = {}
    for name, perm in (("caller", np.arange(g.n_col_eles)), ("shuffled", rng.permutation(g.n_col_eles))):
        s = fc.Solid(0)
        s.set_mesh(fc.FC_HEX8, DISP, g.X, g.node_gid, g.node_owner, g.conn[perm], g.ele_gid[perm])
        s.set_material(NH, 10.0, 0.25)
        n_row, n_col, nnz = s.build_graph()
        for flags in (fc.FC_EVAL_OWNER_COMPUTED, 0):
            K, f = np.zeros(nnz), np.zeros(n_row)
            assert s.evaluate(fc.FC_ACTION_NLNSTIFF, u, f, K, flags=flags | fc.FC_EVAL_ZERO_FIRST) == (0, -1)
            res[name, flags] = (K, f)
        K2, f2 = np.zeros(nnz), np.zeros(n_row)
        s.evaluate(fc.FC_ACTION_NLNSTIFF, u, f2, K2, flags=fc.FC_EVAL_ZERO_FIRST | fc.FC_EVAL_OWNER_COMPUTED)
        oc = res[name, fc.FC_EVAL_OWNER_COMPUTED]
        assert np.array_equal(K2, oc[0]) and np.array_equal(f2, oc[1])  # no atomics: bit-identical reruns
        s.close()
    K0, f0 = res["caller", 0]
    for key, (K, f) in res.items():
        assert np.max(np.abs(K - K0)) <= 1e-13 * np.max(np.abs(K0)), key
        assert np.max(np.abs(f - f0)) <= 1e-13 * np.max(np.abs(K0)), key
    assert fc.capi.l2_sweep_order(g.X, g.conn, 40)[0]


@pytest.mark.gpu
@pytest.mark.parametrize("ct,form,mat,n,nranks", [(0, DISP, NH, (7, 6, 5), 4), (0, FBAR, NH, (6, 5, 4), 2), (0, DISP, SVK, (6, 5, 4), 3),
                                                  (1, DISP, NH, (3, 3, 2), 3), (1, DISP, NH, (3, 2, 2), 2)])
def test_interior_then_boundary_equals_full_evaluation(fc, ct, form, mat, n, nranks):
    """FC_EVAL_INTERIOR evaluates the elements without ghost nodes (ghost displacements poisoned with NaN: they must not be
    read), FC_EVAL_BOUNDARY the rest after the halo import; together they give the full evaluation"""
    for r in range(nranks):
        g = fc.Grid(ct, n, nranks=nranks, rank=r)
        s = fc.Solid(0, r, nranks)
        s.set_mesh(ct, form, g.X, g.node_gid, g.node_owner, g.conn, g.ele_gid)
        s.set_material(mat, 10.0, 0.3)
        n_row, n_col, nnz = s.build_graph()
        u = field(g.X, 0).reshape(-1).copy()
        K0, f0 = np.zeros(nnz), np.zeros(n_row)
        assert s.evaluate(fc.FC_ACTION_NLNSTIFF, u, f0, K0, flags=fc.FC_EVAL_ZERO_FIRST) == (0, -1)
        # device buffers: what the interior half does not write keeps the caller's 9.0
        import torch
        variants = (0, fc.FC_EVAL_OWNER_COMPUTED) if (ct, form, mat) == (0, DISP, NH) else (0,)
        for extra in variants:
            up = u.copy()
            up[n_row:] = np.nan
            up_d, u_d = torch.from_numpy(up).cuda(), torch.from_numpy(u).cuda()
            K1 = torch.full((nnz,), 9.0, dtype=torch.float64, device="cuda")
            f1 = torch.full((n_row,), 9.0, dtype=torch.float64, device="cuda")
            assert s.evaluate(fc.FC_ACTION_NLNSTIFF, up_d, f1, K1, flags=extra | fc.FC_EVAL_ZERO_FIRST | fc.FC_EVAL_INTERIOR) == (0, -1)
            assert bool(torch.isfinite(K1).all()) and bool(torch.isfinite(f1).all())  # the poisoned ghosts were not read
            if n_col > n_row:
                assert float((K1.cpu() - torch.from_numpy(K0)).abs().max()) > 0  # the boundary elements are still missing
            assert s.evaluate(fc.FC_ACTION_NLNSTIFF, u_d, f1, K1, flags=extra | fc.FC_EVAL_BOUNDARY) == (0, -1)
            assert np.max(np.abs(K1.cpu().numpy() - K0)) <= 1e-13 * np.max(np.abs(K0))
            assert np.max(np.abs(f1.cpu().numpy() - f0)) <= 1e-13 * np.max(np.abs(K0))
            # without FC_EVAL_ZERO_FIRST both halves add to what the caller holds (SparseMatrix::assemble semantics)
            K1.fill_(0.5)
            f1.fill_(0.25)
            s.evaluate(fc.FC_ACTION_NLNSTIFF, u_d, f1, K1, flags=extra | fc.FC_EVAL_INTERIOR)
            s.evaluate(fc.FC_ACTION_NLNSTIFF, u_d, f1, K1, flags=extra | fc.FC_EVAL_BOUNDARY)
            assert np.max(np.abs(K1.cpu().numpy() - 0.5 - K0)) <= 1e-13 * max(np.max(np.abs(K0)), 1.0)
            assert np.max(np.abs(f1.cpu().numpy() - 0.25 - f0)) <= 1e-13 * max(np.max(np.abs(K0)), 1.0)
        # kernels that do not follow the launch-slot order refuse the split
        with pytest.raises(fc.FourCError) as ei:
            s.evaluate(fc.FC_ACTION_NLNSTIFF, u_d, f1, K1, flags=fc.FC_EVAL_INTERIOR | fc.FC_EVAL_FORCE_GENERIC)
        assert ei.value.code == fc.FC_ERR_UNSUPPORTED
        s.close()


@pytest.mark.gpu
@pytest.mark.parametrize("ct", [0, 1, 2])
def test_neumann_loads_match_oracle(fc, oracle, ct):
    """fc_solid_neumann_volume / _surface (TYPE Live) against the oracle's restatement of evaluate_neumann
    (4C_solid_3D_ele_neumann_evaluator.cpp:43-115) and SolidSurface neum_live (4C_solid_3D_ele_surface_evaluate.cpp:225-330),
    with per-Gauss-point function factors, host and device outputs, += semantics"""
    import torch
    import structure_driver as sd
    if ct == 2:
        p = sd.Problem(sd.load("solid_ele_tet10_Standard_dynamic"), oracle.node_param_coords)
        X, gid, conn, egid = p.X, p.node_gid, p.conn, p.ele_gid
        ft, faces = p.face_lists[0][1], p.face_lists[0][2]
    else:
        g = fc.Grid(ct, (3, 2, 2))
        X, gid, conn, egid = g.X * np.array([1.0, 1.3, 0.7]), g.node_gid, g.conn, g.ele_gid
        # faces on the x+ side in the reference's quad4 / quad9 node order
        xi = oracle.node_param_coords(ct)
        loc = [a for a in range(len(xi)) if xi[a, 0] == 1.0]
        want = [(-1, -1), (1, -1), (1, 1), (-1, 1), (0, -1), (1, 0), (0, 1), (-1, 0), (0, 0)][:len(loc)]
        order = [next(a for a in loc if (xi[a, 1], xi[a, 2]) == w) for w in want]
        faces = np.array([en[order] for en in conn if np.all(X[en[order], 0] > 1.0 - 1e-12)], np.int32)
        ft = fc.FC_FACE_QUAD4 if ct == 0 else fc.FC_FACE_QUAD9
    s = fc.Solid(0)
    s.set_mesh(ct, DISP, X, gid, np.zeros(len(gid), np.int32), conn, egid)
    s.set_material(NH, 10.0, 0.25)
    rng = np.random.default_rng(2)
    n_dof = 3 * len(gid)
    # volume
    els = np.arange(len(conn), dtype=np.int32)[::2]
    xg = s.neumann_volume_gp_coords(els)
    fe_ref, xg_ref = oracle.volume_neumann(ct, X[conn[els]], [1, 0, 1], [2.0, 9.0, -0.5], want_coords=True)
    assert np.max(np.abs(xg - xg_ref)) <= 1e-14
    fac = 1.0 + rng.random(xg.shape)
    fe_ref = oracle.volume_neumann(ct, X[conn[els]], [1, 0, 1], [2.0, 9.0, -0.5], fac)
    f_ref = np.full(n_dof, 0.125)
    np.add.at(f_ref.reshape(-1, 3), conn[els], fe_ref)
    f = np.full(n_dof, 0.125)
    s.neumann_volume(els, [1, 0, 1], [2.0, 9.0, -0.5], f, gp_factor=fac)
    assert np.max(np.abs(f - f_ref)) <= 1e-14 * np.max(np.abs(f_ref))
    f_dev = torch.full((n_dof,), 0.125, dtype=torch.float64, device="cuda")
    s.neumann_volume(els, [1, 0, 1], [2.0, 9.0, -0.5], f_dev, gp_factor=torch.from_numpy(fac).cuda())
    assert np.max(np.abs(f_dev.cpu().numpy() - f_ref)) <= 1e-14 * np.max(np.abs(f_ref))
    # surface
    assert len(faces) > 0
    xs = s.neumann_surface_gp_coords(ft, faces)
    fs_ref, xs_ref = oracle.surface_neumann(ft, X[faces], [1, 1, 1], [1.0, -2.0, 0.3], want_coords=True)
    assert np.max(np.abs(xs - xs_ref)) <= 1e-14
    fac = 1.0 + rng.random(xs.shape)
    fs_ref = oracle.surface_neumann(ft, X[faces], [1, 1, 1], [1.0, -2.0, 0.3], fac)
    f_ref = np.zeros(n_dof)
    np.add.at(f_ref.reshape(-1, 3), faces, fs_ref)
    f = np.zeros(n_dof)
    s.neumann_surface(ft, faces, [1, 1, 1], [1.0, -2.0, 0.3], f, gp_factor=fac)
    assert np.max(np.abs(f - f_ref)) <= 1e-14 * np.max(np.abs(f_ref))
    s.close()
