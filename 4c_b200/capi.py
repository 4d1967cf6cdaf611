"""ctypes binding of libfourc_b200.so (include/fourc_b200.h).

Plumbing only: the library is the product.  Buffers may be numpy arrays (host), torch tensors
(host or CUDA; passed by data_ptr) or raw integer addresses.  There is no fallback of any kind: if the
shared library is missing this module raises, and without a CUDA device Solid() raises.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libfourc_b200.so")

FC_HEX8, FC_HEX27, FC_TET10, FC_HEX20, FC_TET4, FC_WEDGE6, FC_PYRAMID5, FC_HEX18, FC_NURBS27 = 0, 1, 2, 3, 4, 5, 6, 7, 8
FC_FORM_DISPLACEMENT, FC_FORM_FBAR, FC_FORM_LINEAR, FC_FORM_MULF, FC_FORM_EAS_MILD, FC_FORM_EAS_FULL, FC_FORM_MULF_FBAR, FC_FORM_EAS_SHELL, FC_FORM_SHELL_ANS, FC_FORM_SHELL_EAS_ANS, FC_FORM_EAS_MILD_LINEAR, FC_FORM_EAS_FULL_LINEAR = 0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11
FC_STRAIN_GL, FC_STRAIN_EA, FC_STRAIN_LOG = 0, 1, 2
FC_MAT_STVENANTKIRCHHOFF, FC_MAT_ELASTHYPER_COUPNEOHOOKE = 0, 1
FC_SUMMAND_COUPNEOHOOKE, FC_SUMMAND_COUPSVK, FC_SUMMAND_COUPANISOEXPO = 0, 1, 2
FC_ACTION_NLNSTIFF, FC_ACTION_INTERNALFORCE = 0, 1
FC_EVAL_ZERO_FIRST, FC_EVAL_FORCE_GENERIC, FC_EVAL_CLOSED_FORM_FMA, FC_EVAL_OWNER_COMPUTED, FC_EVAL_DMMA_TILES = 1, 2, 4, 8, 16
FC_EVAL_INTERIOR, FC_EVAL_BOUNDARY = 32, 64
FC_STRESS_2PK, FC_STRESS_CAUCHY = 0, 1
FC_FACE_QUAD4, FC_FACE_QUAD9, FC_FACE_TRI6, FC_FACE_QUAD8, FC_FACE_TRI3, FC_FACE_QUAD6 = 0, 1, 2, 3, 4, 5
FACE_NEN = {FC_FACE_QUAD4: 4, FC_FACE_QUAD9: 9, FC_FACE_TRI6: 6, FC_FACE_QUAD8: 8, FC_FACE_TRI3: 3, FC_FACE_QUAD6: 6}
FACE_NGP = {FC_FACE_QUAD4: 4, FC_FACE_QUAD9: 9, FC_FACE_TRI6: 6, FC_FACE_QUAD8: 9, FC_FACE_TRI3: 3, FC_FACE_QUAD6: 6}
FC_OK, FC_ERR_INVALID, FC_ERR_NO_DEVICE, FC_ERR_CUDA = 0, -1, -2, -3
FC_ERR_NEGATIVE_JACOBIAN, FC_ERR_GRAPH, FC_ERR_UNSUPPORTED = -4, -5, -6
FC_ERR_SOLVER, FC_ERR_SOLVER_NOT_CONVERGED = -7, -8
FC_SOLVER_CG, FC_SOLVER_BICGSTAB, FC_SOLVER_CG_MG = 0, 1, 2
NEN = {FC_HEX8: 8, FC_HEX27: 27, FC_TET10: 10, FC_HEX20: 20, FC_TET4: 4, FC_WEDGE6: 6, FC_PYRAMID5: 5, FC_HEX18: 18, FC_NURBS27: 27}
NGP = {FC_HEX8: 8, FC_HEX27: 27, FC_TET10: 4, FC_HEX20: 27, FC_TET4: 1, FC_WEDGE6: 6, FC_PYRAMID5: 8, FC_HEX18: 18, FC_NURBS27: 27}

EXPORTS = [
    "fc_version", "fc_last_error", "fc_solid_create", "fc_solid_destroy", "fc_solid_set_stream",
    "fc_solid_set_mesh", "fc_solid_set_material", "fc_solid_set_material_elasthyper", "fc_solid_set_fibers", "fc_solid_set_nodal_fibers", "fc_solid_update_prestress", "fc_solid_recover", "fc_solid_eas_history_doubles", "fc_solid_eas_get_history", "fc_solid_eas_set_history", "fc_solid_evaluate_strain", "fc_solid_mulf_get_history", "fc_solid_mulf_set_history", "fc_solid_set_graph", "fc_solid_build_graph",
    "fc_solid_graph_sizes", "fc_solid_get_graph", "fc_solid_set_dirichlet_rows", "fc_solid_evaluate",
    "fc_solid_error_state", "fc_solid_error_allreduce", "fc_solid_zero", "fc_solid_evaluate_mass", "fc_solid_evaluate_stress", "fc_solid_evaluate_stress_noalloc", "fc_solid_neumann_surface", "fc_solid_neumann_volume",
    "fc_solid_neumann_surface_gp_coords", "fc_solid_neumann_volume_gp_coords", "fc_solid_evaluate_elements",
    "fc_solid_set_nurbs", "fc_solid_neumann_surface_nurbs", "fc_solid_neumann_surface_nurbs_gp_coords", "fc_solid_neumann_surface_orthopressure", "fc_solid_neumann_surface_orthopressure_linearization",
    "fc_solid_set_element_kernel_flags", "fc_solid_halo_set_send", "fc_solid_halo_pack", "fc_solid_halo_set_counts", "fc_solid_halo_exchange",
    "fc_solid_last_kernel_ms", "fc_solid_kernel_ms_history", "fc_solid_kernel_launch_count", "fc_grid_create", "fc_grid_destroy",
    "fc_grid_sizes", "fc_grid_subdivisions", "fc_grid_get_nodes", "fc_grid_get_elements",
    "fc_grid_halo_counts", "fc_grid_halo_send_nodes", "fc_probe_fp64_fma", "fc_probe_fp64_mma",
    "fc_solid_set_comm", "fc_solid_set_dirichlet_toggle", "fc_solid_apply_dirichlet", "fc_solid_apply_dirichlet_to_system", "fc_solid_spmv", "fc_solid_linear_solve", "fc_solid_mg_setup", "fc_solid_mg_update", "fc_solid_mg_set_sweeps", "fc_solid_mg_set_precision", "fc_solid_mg_set_damping", "fc_solid_mg_level_info",
    "fc_l2_sweep_order", "fc_brick_plan_create", "fc_brick_plan_sizes", "fc_brick_plan_get", "fc_brick_plan_destroy",
]


class FourCError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"fourc_b200 error {code}: {msg}")
        self.code = code


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(make -C 4c_b200/csrc).  There is no CPU fallback.")
        _lib = C.CDLL(LIB_PATH)
        _lib.fc_version.restype = C.c_char_p
        _lib.fc_last_error.restype = C.c_char_p
        _lib.fc_last_error.argtypes = [C.c_void_p]
        _lib.fc_solid_destroy.restype = None
        _lib.fc_solid_destroy.argtypes = [C.c_void_p]
        _lib.fc_grid_destroy.restype = None
        _lib.fc_grid_destroy.argtypes = [C.c_void_p]
    return _lib


def _ptr(x):
    """address of a numpy array / torch tensor / int / None as c_void_p"""
    if x is None:
        return C.c_void_p(None)
    if isinstance(x, int):
        return C.c_void_p(x)
    if isinstance(x, np.ndarray):
        assert x.flags["C_CONTIGUOUS"]
        return C.c_void_p(x.ctypes.data)
    if hasattr(x, "data_ptr"):
        assert x.is_contiguous()
        return C.c_void_p(x.data_ptr())
    raise TypeError(type(x))


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def probe_fp64(device=0):
    """(FMA TFLOP/s, tensor-core TFLOP/s) measured with the library's probe kernels"""
    a, b = C.c_double(), C.c_double()
    for f, o in ((lib().fc_probe_fp64_fma, a), (lib().fc_probe_fp64_mma, b)):
        rc = f(C.c_int(device), C.byref(o))
        if rc != 0:
            raise FourCError(rc, "fp64 probe")
    return a.value, b.value


def brick_plan(X, conn, n_row_nodes, max_row_nodes=27, chunk=0):
    """brick plan of the owner-computed hex8 assembly as plain arrays (fc_brick_plan_*; host integer code, no device needed):
    dict with n_units, n_interior_units, chunk, n_element_instances, unit_planes, unit_node_off, unit_ele_off, node_tab,
    ele_tab, rowidx_tab [entries, 8] (bytes); None if the mesh is not a lattice of hex cells"""
    L = lib()
    X, conn = _f64(X), _i32(conn)
    h = C.c_void_p()
    rc = L.fc_brick_plan_create(C.c_int(X.size // 3), _ptr(X), C.c_int(conn.size // 8), _ptr(conn), C.c_int(n_row_nodes),
                                C.c_int(max_row_nodes), C.c_int(chunk), C.byref(h))
    if rc == FC_ERR_GRAPH:
        return None
    if rc != 0:
        raise FourCError(rc, "fc_brick_plan_create")
    nu, ni, ch, nn, ne, inst = C.c_int(), C.c_int(), C.c_int(), C.c_int64(), C.c_int64(), C.c_int64()
    L.fc_brick_plan_sizes(h, C.byref(nu), C.byref(ni), C.byref(ch), C.byref(nn), C.byref(ne), C.byref(inst))
    out = dict(n_units=nu.value, n_interior_units=ni.value, chunk=ch.value, n_element_instances=inst.value,
               unit_planes=np.zeros(nu.value, np.int32), unit_node_off=np.zeros(nu.value, np.int64), unit_ele_off=np.zeros(nu.value, np.int64),
               node_tab=np.zeros(nn.value, np.int32), ele_tab=np.zeros(ne.value, np.int32), rowidx_tab=np.zeros(ne.value, np.uint64))
    L.fc_brick_plan_get(h, _ptr(out["unit_planes"]), _ptr(out["unit_node_off"]), _ptr(out["unit_ele_off"]), _ptr(out["node_tab"]),
                        _ptr(out["ele_tab"]), _ptr(out["rowidx_tab"]))
    L.fc_brick_plan_destroy.restype = None
    L.fc_brick_plan_destroy(h)
    out["rowidx_tab"] = out["rowidx_tab"].view(np.uint8).reshape(-1, 8)  # little endian: byte a = local node a
    return out


def l2_sweep_order(X, conn, layer_elems):
    """(changed, perm): element order of the hex8 assembled-path kernels on large meshes (fc_l2_sweep_order)"""
    X, conn = _f64(X), _i32(conn)
    perm = np.zeros(conn.shape[0], np.int32)
    rc = lib().fc_l2_sweep_order(C.c_int(X.size // 3), _ptr(X), C.c_int(conn.shape[0]), C.c_int(conn.shape[1]), _ptr(conn), C.c_int64(layer_elems), _ptr(perm))
    if rc < 0:
        raise FourCError(rc, "fc_l2_sweep_order")
    return bool(rc), perm


class _NcclUniqueId(C.Structure):
    _fields_ = [("internal", C.c_byte * 128)]


def nccl_comm_init(rank, world, device_index):
    """An ncclComm_t (as int) for fc_solid_halo_exchange / fc_solid_set_comm, created with the NCCL that torch ships
    (the library resolves the same symbols at run time).  The unique id travels from rank 0 through torch.distributed,
    which must be initialised.  Plumbing for bench.py and the examples; a 4C build hands over its own communicator."""
    import torch
    import torch.distributed as dist
    nccl = None
    for cand in (os.path.join(os.path.dirname(torch.__file__), "..", "nvidia", "nccl", "lib", "libnccl.so.2"), "libnccl.so.2"):
        try:
            nccl = C.CDLL(cand, mode=C.RTLD_GLOBAL)
            break
        except OSError:
            continue
    if nccl is None:
        raise ImportError("libnccl.so.2 not found")
    uid = _NcclUniqueId()
    if rank == 0:
        rc = nccl.ncclGetUniqueId(C.byref(uid))
        if rc != 0:
            raise FourCError(rc, "ncclGetUniqueId")
    t = torch.tensor(list(bytes(uid)), dtype=torch.uint8, device=torch.device("cuda", device_index))
    dist.broadcast(t, 0)
    C.memmove(C.byref(uid), bytes(t.cpu().tolist()), 128)
    comm = C.c_void_p()
    nccl.ncclCommInitRank.argtypes = [C.POINTER(C.c_void_p), C.c_int, _NcclUniqueId, C.c_int]
    torch.cuda.set_device(device_index)
    rc = nccl.ncclCommInitRank(C.byref(comm), C.c_int(world), uid, C.c_int(rank))
    if rc != 0:
        raise FourCError(rc, "ncclCommInitRank")
    return comm.value


class Grid:
    """Structured cuboid + reference decomposition for one rank (fc_grid_*)."""

    def __init__(self, celltype, n, lo=(0.0, 0.0, 0.0), hi=(1.0, 1.0, 1.0), nranks=1, rank=0, node_offset=0):
        L = lib()
        h = C.c_void_p()
        lo, hi = _f64(lo), _f64(hi)
        rc = L.fc_grid_create(C.byref(h), C.c_int(celltype), C.c_int(n[0]), C.c_int(n[1]), C.c_int(n[2]),
                              _ptr(lo), _ptr(hi), C.c_int(nranks), C.c_int(rank), C.c_int(node_offset))
        if rc != 0:
            raise FourCError(rc, "fc_grid_create")
        self.h, self.celltype, self.n, self.nranks, self.rank = h, celltype, tuple(n), nranks, rank
        s = np.zeros(6, np.int64)
        L.fc_grid_sizes(h, _ptr(s))
        self.n_row_nodes, self.n_col_nodes, self.n_row_eles, self.n_col_eles = (int(v) for v in s[:4])
        self.min_node_gid, self.nen = int(s[4]), int(s[5])
        self.node_gid = np.zeros(self.n_col_nodes, np.int32)
        self.node_owner = np.zeros(self.n_col_nodes, np.int32)
        self.X = np.zeros((self.n_col_nodes, 3))
        L.fc_grid_get_nodes(h, _ptr(self.node_gid), _ptr(self.node_owner), _ptr(self.X))
        self.ele_gid = np.zeros(self.n_col_eles, np.int32)
        self.ele_owner = np.zeros(self.n_col_eles, np.int32)
        self.conn = np.zeros((self.n_col_eles, self.nen), np.int32)
        L.fc_grid_get_elements(h, _ptr(self.ele_gid), _ptr(self.ele_owner), _ptr(self.conn))
        sub = np.zeros(3, np.int32)
        L.fc_grid_subdivisions(h, _ptr(sub))
        self.subdivisions = tuple(int(v) for v in sub)
        self.send_count = np.zeros(nranks, np.int32)
        self.recv_count = np.zeros(nranks, np.int32)
        L.fc_grid_halo_counts(h, _ptr(self.send_count), _ptr(self.recv_count))
        self.send_nodes = np.zeros(int(self.send_count.sum()), np.int32)
        if len(self.send_nodes):
            L.fc_grid_halo_send_nodes(h, _ptr(self.send_nodes))

    def dof_gid(self, node_gid):
        """4C_fem_dofset.cpp:344-352: 3 (gid - min gid) + j"""
        g = (np.asarray(node_gid, np.int64) - self.min_node_gid) * 3
        return (g[:, None] + np.arange(3)[None, :]).reshape(-1)

    def __del__(self):
        try:
            lib().fc_grid_destroy(self.h)
        except Exception:
            pass


class Solid:
    """One rank's context (fc_solid_*)."""

    def __init__(self, device=0, rank=0, nranks=1):
        L = lib()
        self.h = C.c_void_p()
        rc = L.fc_solid_create(C.byref(self.h), C.c_int(device), C.c_int(rank), C.c_int(nranks))
        if rc != 0:
            self.h = C.c_void_p()
            raise FourCError(rc, L.fc_last_error(None).decode() or "fc_solid_create failed")
        self.rank, self.nranks, self.device = rank, nranks, device

    def _chk(self, rc):
        if rc != 0:
            raise FourCError(rc, lib().fc_last_error(self.h).decode())

    def set_stream(self, stream_handle):
        self._chk(lib().fc_solid_set_stream(self.h, C.c_void_p(stream_handle)))

    def set_mesh(self, celltype, formulation, X, node_gid, node_owner, conn, ele_gid):
        X, node_gid, node_owner, conn, ele_gid = _f64(X), _i32(node_gid), _i32(node_owner), _i32(conn), _i32(ele_gid)
        self.celltype, self.nen, self.formulation = celltype, NEN[celltype], formulation
        self.n_col_nodes, self.n_ele = len(node_gid), len(ele_gid)
        self.n_col_dof = 3 * self.n_col_nodes
        self.n_row_dof = 3 * int(np.sum(node_owner == self.rank))
        self._chk(lib().fc_solid_set_mesh(self.h, C.c_int(celltype), C.c_int(formulation), C.c_int(self.n_col_nodes),
                                          _ptr(X), _ptr(node_gid), _ptr(node_owner), C.c_int(self.n_ele), _ptr(conn),
                                          _ptr(ele_gid)))

    def set_nurbs(self, ele_knots, node_weights):
        """FC_NURBS27: knot spans (n_col_ele, 3, 6) and control point weights (n_col_nodes,); must follow set_mesh"""
        kn, w = _f64(ele_knots).reshape(-1), _f64(node_weights)
        assert kn.size == 18 * self.n_ele and w.size == self.n_col_nodes
        self._chk(lib().fc_solid_set_nurbs(self.h, _ptr(kn), _ptr(w)))

    def set_material(self, kind, young, nue):
        p = _f64([young, nue])
        self._chk(lib().fc_solid_set_material(self.h, C.c_int(kind), _ptr(p), C.c_int(2)))

    def set_material_elasthyper(self, summands, fibers=None):
        """MAT_ElastHyper with a summand list [(kind, [4 parameters])]; fibers: (n_ele, 3) FIBER1 per element or None"""
        kinds, par = _i32([k for k, _ in summands]), _f64([list(q) + [0.0] * (4 - len(q)) for _, q in summands])
        self._chk(lib().fc_solid_set_material_elasthyper(self.h, C.c_int(len(kinds)), _ptr(kinds), _ptr(par)))
        fb = None if fibers is None else _f64(fibers)
        self._chk(lib().fc_solid_set_fibers(self.h, _ptr(fb)))

    def set_nodal_fibers(self, node_fibers):
        """(n_col_nodes, 3) FIBER1 of the fibre nodes: interpolated to the Gauss points by the library; None clears"""
        fb = None if node_fibers is None else _f64(node_fibers)
        self._chk(lib().fc_solid_set_nodal_fibers(self.h, _ptr(fb)))

    def update_prestress(self, u_col):
        """FC_FORM_MULF: fold the displacements of the finished step into the Gauss-point history (the caller resets them)"""
        self._chk(lib().fc_solid_update_prestress(self.h, _ptr(u_col if hasattr(u_col, "data_ptr") else _f64(u_col))))

    def recover(self, du_col, step_length=1.0):
        """struct_calc_recover (EAS): alpha += step_length * alpha_inc(du_col); a no-op for formulations without condensed variables"""
        self._chk(lib().fc_solid_recover(self.h, _ptr(du_col if hasattr(du_col, "data_ptr") else _f64(du_col)), C.c_double(step_length)))

    def eas_history(self):
        """(n_col_ele, 3 n + n^2 + 24 n + 1): alpha_inc | alpha | s | Kaa^-1 | Kda | old_step_length of every element"""
        n = C.c_int()
        self._chk(lib().fc_solid_eas_history_doubles(self.h, C.byref(n)))
        h = np.zeros((self.n_ele, n.value))
        self._chk(lib().fc_solid_eas_get_history(self.h, _ptr(h)))
        return h

    def set_eas_history(self, history):
        self._chk(lib().fc_solid_eas_set_history(self.h, _ptr(_f64(history))))

    def mulf_history(self):
        """(n_col_ele, ngp, 18) inverse Jacobian | deformation gradient of every Gauss point"""
        h = np.zeros((self.n_ele, NGP[self.celltype] + (1 if self.formulation == FC_FORM_MULF_FBAR else 0), 18))
        self._chk(lib().fc_solid_mulf_get_history(self.h, _ptr(h)))
        return h

    def set_mulf_history(self, history):
        h = _f64(history)
        assert h.size == self.n_ele * (NGP[self.celltype] + (1 if self.formulation == FC_FORM_MULF_FBAR else 0)) * 18
        self._chk(lib().fc_solid_mulf_set_history(self.h, _ptr(h)))

    def build_graph(self):
        self._chk(lib().fc_solid_build_graph(self.h))
        return self.graph_sizes()

    def set_graph(self, rowptr, colind):
        rowptr, colind = np.ascontiguousarray(rowptr, np.int64), _i32(colind)
        self._chk(lib().fc_solid_set_graph(self.h, C.c_int(len(rowptr) - 1), _ptr(rowptr), _ptr(colind),
                                           C.c_int(self.n_col_dof)))
        return self.graph_sizes()

    def graph_sizes(self):
        a, b, n = C.c_int(), C.c_int(), C.c_int64()
        self._chk(lib().fc_solid_graph_sizes(self.h, C.byref(a), C.byref(b), C.byref(n)))
        self.n_row_dof, self.n_col_dof, self.nnz = a.value, b.value, n.value
        return a.value, b.value, n.value

    def get_graph(self):
        rowptr = np.zeros(self.n_row_dof + 1, np.int64)
        colind = np.zeros(self.nnz, np.int32)
        self._chk(lib().fc_solid_get_graph(self.h, _ptr(rowptr), _ptr(colind)))
        return rowptr, colind

    def set_dirichlet_rows(self, rows):
        rows = _i32(rows)
        self._chk(lib().fc_solid_set_dirichlet_rows(self.h, C.c_int(len(rows)), _ptr(rows)))

    # ---- Dirichlet application + linear solve on the device (all buffers are device tensors) ----
    def set_dirichlet_toggle(self, rows):
        rows = _i32(rows)
        self._chk(lib().fc_solid_set_dirichlet_toggle(self.h, C.c_int(len(rows)), _ptr(rows)))

    def apply_dirichlet(self, K_values, diagonalblock=True):
        self._chk(lib().fc_solid_apply_dirichlet(self.h, _ptr(K_values), C.c_int(1 if diagonalblock else 0)))

    def apply_dirichlet_to_system(self, x=None, b=None, dbcval=None):
        self._chk(lib().fc_solid_apply_dirichlet_to_system(self.h, _ptr(x), _ptr(b), _ptr(dbcval)))

    def spmv(self, K_values, x_col, y_row):
        self._chk(lib().fc_solid_spmv(self.h, _ptr(K_values), _ptr(x_col), _ptr(y_row)))

    def mg_setup(self, max_levels=8):
        """geometric multigrid hierarchy for FC_SOLVER_CG_MG; returns the number of levels"""
        n = C.c_int(0)
        self._chk(lib().fc_solid_mg_setup(self.h, C.c_int(max_levels), C.byref(n)))
        return n.value

    def mg_set_sweeps(self, nu):
        self._chk(lib().fc_solid_mg_set_sweeps(self.h, C.c_int(nu)))

    def mg_set_precision(self, single_precision_cycle):
        self._chk(lib().fc_solid_mg_set_precision(self.h, C.c_int(1 if single_precision_cycle else 0)))

    def mg_set_damping(self, factor):
        self._chk(lib().fc_solid_mg_set_damping(self.h, C.c_double(factor)))

    def mg_level_info(self, level):
        """(nodes, Jacobi damping) of one multigrid level, 0 = finest"""
        n, w = C.c_int(0), C.c_double(0.0)
        self._chk(lib().fc_solid_mg_level_info(self.h, C.c_int(level), C.byref(n), C.byref(w)))
        return n.value, w.value

    def mg_update(self, u_col_dev):
        self._chk(lib().fc_solid_mg_update(self.h, _ptr(u_col_dev)))

    def linear_solve(self, K_values, b, x, method=FC_SOLVER_CG, rel_tol=1e-10, max_iter=10000, check=True):
        """K x = b; x holds the initial guess.  Returns (iterations, relative residual)."""
        it, rr = C.c_int(0), C.c_double(0.0)
        rc = lib().fc_solid_linear_solve(self.h, C.c_int(method), _ptr(K_values), _ptr(b), _ptr(x), C.c_double(rel_tol), C.c_int(max_iter),
                                         C.byref(it), C.byref(rr))
        if rc != FC_ERR_SOLVER_NOT_CONVERGED or check:
            self._chk(rc)
        return it.value, rr.value

    def evaluate(self, action, u_col, f_row=None, K_values=None, flags=0, want_errors=True):
        """returns (ele_error_count, first_bad_gid) or None when want_errors is False (asynchronous)."""
        cnt, gid = C.c_int(0), C.c_int(-1)
        rc = lib().fc_solid_evaluate(self.h, C.c_int(action), C.c_int(flags), _ptr(u_col), _ptr(f_row), _ptr(K_values),
                                     C.byref(cnt) if want_errors else None, C.byref(gid) if want_errors else None)
        if rc == FC_ERR_NEGATIVE_JACOBIAN and want_errors:
            return cnt.value, gid.value
        self._chk(rc)
        return (cnt.value, gid.value) if want_errors else None

    def error_state(self):
        cnt, gid = C.c_int(0), C.c_int(-1)
        rc = lib().fc_solid_error_state(self.h, C.byref(cnt), C.byref(gid))
        if rc not in (FC_OK, FC_ERR_NEGATIVE_JACOBIAN):
            self._chk(rc)
        return cnt.value, gid.value

    def error_allreduce(self):
        """eval_error_check: 1 if any rank of the communicator (set_comm) had an element failure in its last evaluation"""
        flag = C.c_int(0)
        self._chk(lib().fc_solid_error_allreduce(self.h, C.byref(flag)))
        return flag.value

    def set_comm(self, nccl_comm):
        self._chk(lib().fc_solid_set_comm(self.h, C.c_void_p(nccl_comm)))

    def halo_set_counts(self, send_dof_count, recv_dof_count):
        s_, r_ = _i32(send_dof_count), _i32(recv_dof_count)
        self._chk(lib().fc_solid_halo_set_counts(self.h, _ptr(s_), _ptr(r_)))

    def halo_exchange(self, nccl_comm, u_col_dev):
        self._chk(lib().fc_solid_halo_exchange(self.h, C.c_void_p(nccl_comm), _ptr(u_col_dev)))

    def zero(self, f_row=None, K_values=None):
        self._chk(lib().fc_solid_zero(self.h, _ptr(f_row), _ptr(K_values)))

    def evaluate_mass(self, density, M_values, flags=FC_EVAL_ZERO_FIRST):
        """consistent mass (struct_calc_nlnstiffmass) into M_values [nnz], layout of the stiffness graph"""
        self._chk(lib().fc_solid_evaluate_mass(self.h, C.c_double(density), C.c_int(flags), _ptr(M_values)))

    # ---- Neumann loads TYPE Live ----
    def neumann_surface(self, face_type, face_nodes, onoff, val, f_ext, gp_factor=None):
        fn = _i32(face_nodes).reshape(-1, FACE_NEN[face_type])
        oo, vv = _i32(onoff[:3]), _f64(val[:3])
        gf = None if gp_factor is None else (_f64(gp_factor) if isinstance(gp_factor, np.ndarray) else gp_factor)
        self._chk(lib().fc_solid_neumann_surface(self.h, C.c_int(face_type), C.c_int(len(fn)), _ptr(fn), _ptr(oo), _ptr(vv), _ptr(gf), _ptr(f_ext)))

    def neumann_volume(self, ele_lids, onoff, val, f_ext, gp_factor=None):
        el = _i32(ele_lids)
        oo, vv = _i32(onoff[:3]), _f64(val[:3])
        gf = None if gp_factor is None else (_f64(gp_factor) if isinstance(gp_factor, np.ndarray) else gp_factor)
        self._chk(lib().fc_solid_neumann_volume(self.h, C.c_int(len(el)), _ptr(el), _ptr(oo), _ptr(vv), _ptr(gf), _ptr(f_ext)))

    def neumann_surface_gp_coords(self, face_type, face_nodes):
        fn = _i32(face_nodes).reshape(-1, FACE_NEN[face_type])
        ngp = FACE_NGP[face_type]
        out = np.zeros((len(fn), ngp, 3))
        self._chk(lib().fc_solid_neumann_surface_gp_coords(self.h, C.c_int(face_type), C.c_int(len(fn)), _ptr(fn), _ptr(out)))
        return out

    def neumann_surface_orthopressure(self, face_type, face_nodes, u_col, pressure, f_ext, gp_factor=None):
        """TYPE (pseudo_)orthopressure: pressure along the normal of X + u_col (None = material configuration), force only"""
        fn = _i32(face_nodes).reshape(-1, FACE_NEN[face_type])
        uu = None if u_col is None else (u_col if hasattr(u_col, "data_ptr") else _f64(u_col))
        gf = None if gp_factor is None else (_f64(gp_factor) if isinstance(gp_factor, np.ndarray) else gp_factor)
        self._chk(lib().fc_solid_neumann_surface_orthopressure(self.h, C.c_int(face_type), C.c_int(len(fn)), _ptr(fn), _ptr(uu),
                                                              C.c_double(pressure), _ptr(gf), _ptr(f_ext)))

    def neumann_surface_orthopressure_linearization(self, face_type, face_nodes, u_col, pressure, K_values, gp_factor=None):
        """LOADLIN yes: adds the load stiffness of the orthopressure load into K_values [nnz]"""
        fn = _i32(face_nodes).reshape(-1, FACE_NEN[face_type])
        uu = None if u_col is None else (u_col if hasattr(u_col, "data_ptr") else _f64(u_col))
        gf = None if gp_factor is None else (_f64(gp_factor) if isinstance(gp_factor, np.ndarray) else gp_factor)
        self._chk(lib().fc_solid_neumann_surface_orthopressure_linearization(self.h, C.c_int(face_type), C.c_int(len(fn)), _ptr(fn), _ptr(uu),
                                                                            C.c_double(pressure), _ptr(gf), _ptr(K_values)))

    def neumann_surface_nurbs(self, parent_ele, surface_id, onoff, val, f_ext, gp_factor=None):
        """faces of nurbs27 elements, named by parent column element and the parent's surface number 0..5"""
        pe, sid = _i32(parent_ele), _i32(surface_id)
        oo, vv = _i32(onoff[:3]), _f64(val[:3])
        gf = None if gp_factor is None else (_f64(gp_factor) if isinstance(gp_factor, np.ndarray) else gp_factor)
        self._chk(lib().fc_solid_neumann_surface_nurbs(self.h, C.c_int(len(pe)), _ptr(pe), _ptr(sid), _ptr(oo), _ptr(vv), _ptr(gf), _ptr(f_ext)))

    def neumann_surface_nurbs_gp_coords(self, parent_ele, surface_id):
        pe, sid = _i32(parent_ele), _i32(surface_id)
        out = np.zeros((len(pe), 9, 3))
        self._chk(lib().fc_solid_neumann_surface_nurbs_gp_coords(self.h, C.c_int(len(pe)), _ptr(pe), _ptr(sid), _ptr(out)))
        return out

    def neumann_volume_gp_coords(self, ele_lids):
        el = _i32(ele_lids)
        ngp = NGP[self.celltype]
        out = np.zeros((len(el), ngp, 3))
        self._chk(lib().fc_solid_neumann_volume_gp_coords(self.h, C.c_int(len(el)), _ptr(el), _ptr(out)))
        return out

    def evaluate_stress(self, u_col, stress_type=1, want_gp=True, want_nodal=True):
        """struct_calc_stress: (gp_stress [n_ele, ngp, 6], nodal_stress [n_row_nodes, 6]); 0 = 2PK, 1 = Cauchy"""
        ngp = NGP[self.celltype]
        gp = np.zeros((self.n_ele, ngp, 6)) if want_gp else None
        nodal = np.zeros((self.n_row_dof // 3, 6)) if want_nodal else None
        self._chk(lib().fc_solid_evaluate_stress(self.h, _ptr(u_col), C.c_int(stress_type), _ptr(gp), _ptr(nodal)))
        return gp, nodal

    def evaluate_strain(self, u_col, strain_type=0, want_gp=True, want_nodal=True):
        """(gp_strain [n_ele, ngp, 6], nodal_strain [n_row_nodes, 6]); 0 = Green-Lagrange, 1 = Euler-Almansi, 2 = logarithmic"""
        ngp = NGP[self.celltype]
        gp = np.zeros((self.n_ele, ngp, 6)) if want_gp else None
        nodal = np.zeros((self.n_row_dof // 3, 6)) if want_nodal else None
        self._chk(lib().fc_solid_evaluate_strain(self.h, _ptr(u_col), C.c_int(strain_type), _ptr(gp), _ptr(nodal)))
        return gp, nodal

    def evaluate_stress_noalloc(self, u_col_dev, nodal_dev, stress_type=1):
        """fc_solid_evaluate_stress_noalloc: nodal stresses into the device tensor nodal_dev [n_row_nodes, 6]; allocates
        nothing, does not synchronise (fixed scratch inside the library, see include/fourc_b200.h for the limits)"""
        self._chk(lib().fc_solid_evaluate_stress_noalloc(self.h, _ptr(u_col_dev), C.c_int(stress_type), _ptr(nodal_dev)))

    def evaluate_elements(self, u_col, want_K=True, flags=0):
        nd = 3 * self.nen
        fe = np.zeros((self.n_ele, nd))
        Ke = np.zeros((self.n_ele, nd, nd)) if want_K else None
        self._chk(lib().fc_solid_set_element_kernel_flags(self.h, C.c_int(flags)))
        self._chk(lib().fc_solid_evaluate_elements(self.h, _ptr(u_col), _ptr(fe), _ptr(Ke)))
        if want_K:
            Ke = np.ascontiguousarray(Ke.transpose(0, 2, 1))  # column-major per element -> [e, r, c]
        return fe, Ke

    def halo_set_send(self, send_row_lids):
        s = _i32(send_row_lids)
        self._chk(lib().fc_solid_halo_set_send(self.h, C.c_int(len(s)), _ptr(s)))

    def halo_pack(self, u_dev, sendbuf_dev):
        self._chk(lib().fc_solid_halo_pack(self.h, _ptr(u_dev), _ptr(sendbuf_dev)))

    def last_kernel_ms(self):
        a, b = C.c_float(), C.c_float()
        self._chk(lib().fc_solid_last_kernel_ms(self.h, C.byref(a), C.byref(b)))
        return a.value, b.value

    def kernel_ms_history(self, n):
        """(zero_ms [n], element_ms [n]) of the last n evaluate / spmv calls, oldest first"""
        z, e = np.zeros(n, np.float32), np.zeros(n, np.float32)
        self._chk(lib().fc_solid_kernel_ms_history(self.h, C.c_int(n), _ptr(z), _ptr(e)))
        return z, e

    def launch_count(self):
        n = C.c_int64()
        self._chk(lib().fc_solid_kernel_launch_count(self.h, C.byref(n)))
        return n.value

    def close(self):
        if self.h:
            lib().fc_solid_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
