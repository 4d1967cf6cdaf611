"""ctypes binding of the CPU oracle (oracle/liboracle.so) — TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module.  The product package (4c_b200/) never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")

HEX8, HEX27, TET10, HEX20, TET4, WEDGE6, PYRAMID5, HEX18, NURBS27 = 0, 1, 2, 3, 4, 5, 6, 7, 8
FORM_DISP, FORM_FBAR, FORM_LINEAR, FORM_MULF, FORM_EAS_MILD, FORM_EAS_FULL, FORM_MULF_FBAR, FORM_EAS_SHELL, FORM_SHELL_ANS, FORM_SHELL_EAS_ANS, FORM_EAS_MILD_LINEAR, FORM_EAS_FULL_LINEAR = 0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11
NEAS = {FORM_EAS_MILD: 9, FORM_EAS_FULL: 21, FORM_EAS_SHELL: 7, FORM_EAS_MILD_LINEAR: 9, FORM_EAS_FULL_LINEAR: 21}
OUTPUT_STRAIN = 16  # stress_type = OUTPUT_STRAIN + strain type: the Gauss-point output is that strain (stress-like Voigt)
STRAIN_GL, STRAIN_EA, STRAIN_LOG = 0, 1, 2
MAT_SVK, MAT_NEOHOOKE = 0, 1
ACTION_NLNSTIFF, ACTION_INTERNALFORCE = 0, 1
NEN = {HEX8: 8, HEX27: 27, TET10: 10, HEX20: 20, TET4: 4, WEDGE6: 6, PYRAMID5: 5, HEX18: 18, NURBS27: 27}
NGP = {HEX8: 8, HEX27: 27, TET10: 4, HEX20: 27, TET4: 1, WEDGE6: 6, PYRAMID5: 8, HEX18: 18, NURBS27: 27}

_dp = np.ctypeslib.ndpointer(dtype=np.float64, flags="C_CONTIGUOUS")
_ip = np.ctypeslib.ndpointer(dtype=np.int32, flags="C_CONTIGUOUS")
_lp = np.ctypeslib.ndpointer(dtype=np.int64, flags="C_CONTIGUOUS")


def build():
    """(Re)build liboracle.so (and oracle/_ref when /root/reference is present)."""
    subprocess.run(["make", "-s", "-C", ORACLE_DIR], check=True, stdout=subprocess.DEVNULL)


def _load():
    path = os.path.join(ORACLE_DIR, "liboracle.so")
    if not os.path.exists(path):
        build()
    return _bind(C.CDLL(path))


def _bind(lib):
    lib.oracle_svk_strain_energy.restype = C.c_double
    lib.oracle_voigt_determinant.restype = C.c_double
    lib.oracle_invert3x3.restype = C.c_double
    lib.oracle_determinant3x3.restype = C.c_double
    for name in ("oracle_job_create_cuboid", "oracle_job_create_mesh"):
        getattr(lib, name).restype = C.c_void_p
    return lib


lib = _load()


def use_fast_build():
    """Rebind this module to the TIMING build of the oracle (bench.py's cpu_baseline / --impl reference legs only):
    -O3 -march=native with FMA contraction, compiled on the machine it runs on (one file per CPU model; falls back to
    x86-64-v3 with contraction if the native build fails).  Returns the compiler flags used.  Parity tests never call this."""
    global lib
    import hashlib
    try:
        with open("/proc/cpuinfo") as f:
            flags = next((ln for ln in f if ln.startswith("flags")), "")
    except OSError:
        flags = ""
    tag = hashlib.sha256(flags.encode()).hexdigest()[:10]
    base = "-std=c++20 -O3 -fPIC -Wall -Wno-unused-function"
    for name, fl in ((f"liboracle_fast_{tag}.so", base + " -march=native"), ("liboracle_fast_v3.so", base + " -march=x86-64-v3")):
        path = os.path.join(ORACLE_DIR, name)
        if not os.path.exists(path):
            r = subprocess.run(["make", "-s", "-C", ORACLE_DIR, "fast", f"FAST_OUT={name}", f"FASTFLAGS={fl}"], capture_output=True)
            if r.returncode != 0:
                continue
        lib = _bind(C.CDLL(path))
        return fl
    raise RuntimeError("the timing build of the oracle could not be compiled")


def _d(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def _i(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def gamma_delta(prinv, dPI, ddPII):
    g, d = np.zeros(3), np.zeros(8)
    lib.oracle_calculate_gamma_delta(_p(_d(prinv)), _p(_d(dPI)), _p(_d(ddPII)), _p(g), _p(d))
    return g, d


def material_evaluate(kind, young, nue, F, E):
    S, cmat = np.zeros(6), np.zeros(36)
    F = _d(np.asarray(F).T.reshape(-1)) if F is not None else np.zeros(9)
    lib.oracle_material_evaluate(C.c_int(kind), C.c_double(young), C.c_double(nue), _p(F), _p(_d(E)), _p(S), _p(cmat))
    return S, cmat.reshape(6, 6).T.copy()  # col-major -> [r, c]


def element_evaluate(ct, form, kind, young, nue, X, u, want_K=True, want_M=False, rho=0.0):
    """X, u: (nen, 3).  Returns err, fe (3*nen), Ke (nd, nd) [r, c], Me."""
    nd = 3 * NEN[ct]
    X, u = _d(X).reshape(-1), _d(u).reshape(-1)
    fe = np.zeros(nd)
    Ke = np.zeros(nd * nd) if want_K else None
    Me = np.zeros(nd * nd) if want_M else None
    err = lib.oracle_element_evaluate(C.c_int(ct), C.c_int(form), C.c_int(kind), C.c_double(young),
                                      C.c_double(nue), _p(X), _p(u), _p(fe), _p(Ke), _p(Me), C.c_double(rho))
    Ke = Ke.reshape(nd, nd).T.copy() if want_K else None
    Me = Me.reshape(nd, nd).T.copy() if want_M else None
    return err, fe, Ke, Me


def element_evaluate_elasthyper(ct, form, summands, fiber, X, u, stress_type=1):
    """one element of MAT_ElastHyper with a summand list: err, fe, Ke [r, c], gp_stress (ngp, 6); fiber: (3,) FIBER1 of the
    element, (ngp, 3) fibres at the Gauss points, or None"""
    nd = 3 * NEN[ct]
    kinds, par = _i([k for k, _ in summands]), _d([q for _, q in summands])
    fe, Ke, gp = np.zeros(nd), np.zeros(nd * nd), np.zeros((NGP[ct], 6))
    fib = _d(fiber) if fiber is not None else None
    err = lib.oracle_element_evaluate_elasthyper(C.c_int(ct), C.c_int(form), C.c_int(len(summands)), _p(kinds), _p(par), _p(fib),
                                                 _p(_d(X).reshape(-1)), _p(_d(u).reshape(-1)), _p(fe), _p(Ke), _p(gp), C.c_int(stress_type),
                                                 C.c_int(3 if fib is not None and fib.size > 3 else 0))
    return err, fe, Ke.reshape(nd, nd).T.copy(), gp


def project_fibers_to_gauss_points(ct, nodal):
    """nodal (nen, 3) -> (ngp, 3): fibres interpolated to the Gauss points and normalised"""
    out = np.zeros((NGP[ct], 3))
    lib.oracle_project_fibers_to_gauss_points(C.c_int(ct), _p(_d(nodal)), _p(out))
    return out


def element_gp_stress(ct, form, kind, young, nue, X, u, stress_type=1):
    """Gauss-point stresses (ngp, 6) of one element; stress_type 0 = 2nd Piola-Kirchhoff, 1 = Cauchy"""
    X, u = _d(X), _d(u)
    out = np.zeros((NGP[ct], 6))
    err = lib.oracle_element_gp_stress(C.c_int(ct), C.c_int(form), C.c_int(kind), C.c_double(young), C.c_double(nue), _p(X), _p(u),
                                       C.c_int(stress_type), _p(out))
    return err, out


def eas_history_doubles(neas):
    return lib.oracle_eas_history_doubles(C.c_int(neas))


def element_evaluate_eas(neas, kind, young, nue, X, u, hist, want_K=True, linear=False):
    """one hex8 EAS element with history hist (modified in place: invKaa, Kda, s) -> err, fe (8, 3), Ke (24, 24) [r, c]"""
    X, u = _d(X), _d(u)
    assert hist.dtype == np.float64 and hist.flags["C_CONTIGUOUS"] and hist.size == eas_history_doubles(neas)
    fe, Ke = np.zeros(24), (np.zeros((24, 24)) if want_K else None)
    err = lib.oracle_element_evaluate_eas(C.c_int(neas), C.c_int(kind), C.c_double(young), C.c_double(nue), _p(X), _p(u), _p(hist),
                                          _p(fe), _p(Ke) if want_K else None, C.c_int(1 if linear else 0))
    return err, fe.reshape(8, 3), (np.ascontiguousarray(Ke.T) if want_K else None)


def element_recover_eas(neas, hist, du, step_length=1.0):
    lib.oracle_element_recover_eas(C.c_int(neas), _p(hist), _p(_d(du)), C.c_double(step_length))


QUAD4, QUAD9, TRI6, QUAD8, TRI3, QUAD6, NURBS9 = 0, 1, 2, 3, 4, 5, 6
FACE_NEN = {QUAD4: 4, QUAD9: 9, TRI6: 6, QUAD8: 8, TRI3: 3, QUAD6: 6, NURBS9: 9}
FACE_NGP = {QUAD4: 4, QUAD9: 9, TRI6: 6, QUAD8: 9, TRI3: 3, QUAD6: 6, NURBS9: 9}
# faces of a nurbs27 element and the parent directions of their two knot vectors (fc_oracle_nurbs.hpp)
NURBS27_FACES = [[0, 1, 2, 3, 4, 5, 6, 7, 8], [18, 19, 20, 21, 22, 23, 24, 25, 26], [0, 1, 2, 9, 10, 11, 18, 19, 20],
                 [6, 7, 8, 15, 16, 17, 24, 25, 26], [2, 5, 8, 11, 14, 17, 20, 23, 26], [0, 3, 6, 9, 12, 15, 18, 21, 24]]
NURBS27_FACE_DIRS = [(0, 1), (0, 1), (0, 2), (0, 2), (1, 2), (1, 2)]


def set_nurbs_element(knots, weights):
    """knots (3, 6) + weights (27,) of the nurbs27 element, or (2, 6) + (9,) of the nurbs9 face, that the following
    unit-level calls (shape, element_evaluate, surface_neumann, volume_neumann with one item) use"""
    kn, w = _d(knots).reshape(-1), _d(weights).reshape(-1)
    lib.oracle_set_nurbs_element(_p(kn), C.c_int(kn.size), _p(w), C.c_int(w.size))


def surface_neumann(ft, x, onoff, val, functfac=None, want_coords=False):
    """x: (n, nfn, 3) reference coordinates of the face nodes -> fe (n, nfn, 3) [, gp_coord (n, ngp, 3)]"""
    x = _d(x)
    n, nn = x.shape[0], FACE_NEN[ft]
    fe = np.zeros((n, nn, 3))
    gc = np.zeros((n, FACE_NGP[ft], 3)) if want_coords else None
    ff = _d(functfac) if functfac is not None else None
    lib.oracle_surface_neumann(C.c_int(ft), C.c_int(n), _p(x), _p(_i(onoff[:3])), _p(_d(val[:3])), _p(ff), _p(fe), _p(gc))
    return (fe, gc) if want_coords else fe


def surface_neumann_orthopressure(ft, xc, value, functfac=None):
    """xc: (n, nfn, 3) coordinates of the configuration the load type asks for -> fe (n, nfn, 3)"""
    xc = _d(xc)
    fe = np.zeros((xc.shape[0], FACE_NEN[ft], 3))
    ff = _d(functfac) if functfac is not None else None
    lib.oracle_surface_neumann_orthopressure(C.c_int(ft), C.c_int(xc.shape[0]), _p(xc), C.c_double(value), _p(ff), _p(fe))
    return fe


def surface_neumann_orthopressure_linearization(ft, xc, value, functfac=None):
    """LOADLIN yes: load stiffness (n, 3 nfn, 3 nfn) of the orthopressure load on the configuration xc"""
    xc = _d(xc)
    nd = 3 * FACE_NEN[ft]
    Ke = np.zeros((xc.shape[0], nd, nd))
    ff = _d(functfac) if functfac is not None else None
    lib.oracle_surface_neumann_orthopressure_linearization(C.c_int(ft), C.c_int(xc.shape[0]), _p(xc), C.c_double(value), _p(ff), _p(Ke))
    return Ke


def volume_neumann(ct, X, onoff, val, functfac=None, want_coords=False):
    """X: (n, nen, 3) -> fe (n, nen, 3) [, gp_coord (n, ngp, 3)]"""
    X = _d(X)
    n = X.shape[0]
    fe = np.zeros((n, NEN[ct], 3))
    gc = np.zeros((n, NGP[ct], 3)) if want_coords else None
    ff = _d(functfac) if functfac is not None else None
    lib.oracle_volume_neumann(C.c_int(ct), C.c_int(n), _p(X), _p(_i(onoff[:3])), _p(_d(val[:3])), _p(ff), _p(fe), _p(gc))
    return (fe, gc) if want_coords else fe


def gauss_rule(ct):
    xi, w = np.zeros((NGP[ct], 3)), np.zeros(NGP[ct])
    lib.oracle_gauss_rule(C.c_int(ct), _p(xi), _p(w))
    return xi, w


def shape(ct, xi):
    N = np.zeros(NEN[ct])
    lib.oracle_shape(C.c_int(ct), _p(_d(xi)), _p(N), None)
    return N


def node_param_coords(ct):
    xi = np.zeros((NEN[ct], 3))
    lib.oracle_node_param_coords(C.c_int(ct), _p(xi))
    return xi


def element_evaluate_batch(ct, form, kind, young, nue, X, u, want_K=True):
    """X, u: (n, nen, 3).  Returns nbad, fe (n, nd), Ke (n, nd, nd) indexed [e, r, c]."""
    n = X.shape[0]
    nd = 3 * NEN[ct]
    X, u = _d(X).reshape(-1), _d(u).reshape(-1)
    fe = np.zeros((n, nd))
    Ke = np.zeros((n, nd, nd)) if want_K else None
    bad = lib.oracle_element_evaluate_batch(C.c_int(ct), C.c_int(form), C.c_int(kind), C.c_double(young),
                                            C.c_double(nue), C.c_int(n), _p(X), _p(u), _p(fe), _p(Ke))
    if want_K:
        Ke = np.ascontiguousarray(Ke.transpose(0, 2, 1))
    return bad, fe, Ke


class Job:
    """Mesh + decomposition + per-rank CSR systems held by the oracle."""

    def __init__(self, handle):
        self.h = C.c_void_p(handle)
        self.nranks = lib.oracle_job_nranks(self.h)
        nn, ne = lib.oracle_job_n_nodes(self.h), lib.oracle_job_n_eles(self.h)
        self.node_gid = np.zeros(nn, np.int32)
        self.X = np.zeros((nn, 3))
        self.node_owner = np.zeros(nn, np.int32)
        lib.oracle_job_get_nodes(self.h, _p(self.node_gid), _p(self.X), _p(self.node_owner))
        self.ele_gid = np.zeros(ne, np.int32)
        self.ele_owner = np.zeros(ne, np.int32)
        self._conn_flat = None
        self._ne = ne

    @classmethod
    def cuboid(cls, ct, n, lo=(0, 0, 0), hi=(1, 1, 1), nranks=1, node_offset=0):
        lo, hi = _d(lo), _d(hi)
        h = lib.oracle_job_create_cuboid(C.c_int(ct), C.c_int(n[0]), C.c_int(n[1]), C.c_int(n[2]), _p(lo), _p(hi),
                                         C.c_int(nranks), C.c_int(node_offset))
        j = cls(h)
        j.ct = ct
        j._fetch_eles()
        return j

    @classmethod
    def from_mesh(cls, ct, node_gid, X, ele_gid, conn):
        node_gid, X, ele_gid, conn = _i(node_gid), _d(X), _i(ele_gid), _i(conn)
        h = lib.oracle_job_create_mesh(C.c_int(ct), C.c_int(len(node_gid)), _p(node_gid), _p(X),
                                       C.c_int(len(ele_gid)), _p(ele_gid), _p(conn))
        j = cls(h)
        j.ct = ct
        j._fetch_eles()
        return j

    def set_nurbs(self, ele_knots, node_gid, weights):
        """nurbs27: knot spans (n_ele, 3, 6) in the job's element order, weights parallel to node_gid"""
        kn, g, w = _d(ele_knots).reshape(-1), _i(node_gid), _d(weights)
        assert kn.size == 18 * self._ne and g.size == w.size
        lib.oracle_job_set_nurbs(self.h, _p(kn), C.c_int(g.size), _p(g), _p(w))

    def _fetch_eles(self):
        self.conn = np.zeros((self._ne, NEN[self.ct]), np.int32)
        lib.oracle_job_get_eles(self.h, _p(self.ele_gid), _p(self.conn), _p(self.ele_owner))

    def subdivisions(self):
        s = np.zeros(3, np.int32)
        lib.oracle_job_subdivisions(self.h, _p(s))
        return tuple(int(x) for x in s)

    def set_problem(self, form, kind, young, nue):
        lib.oracle_job_set_problem(self.h, C.c_int(form), C.c_int(kind), C.c_double(young), C.c_double(nue))

    def set_elasthyper(self, form, summands, fibers=None):
        """summands: [(kind, [4 parameters])]; fibers: (n_ele, 3) FIBER1 in the order of self.ele_gid or None"""
        kinds = _i([k for k, _ in summands])
        par = _d([q for _, q in summands])
        lib.oracle_job_set_elasthyper(self.h, C.c_int(form), C.c_int(len(summands)), _p(kinds), _p(par),
                                      _p(_d(fibers)) if fibers is not None else None)

    def set_nodal_fibers(self, node_fibers):
        """node_fibers: (n_nodes, 3) FIBER1 of the FNODEs in the order of self.node_gid; call after set_elasthyper"""
        lib.oracle_job_set_nodal_fibers(self.h, _p(_d(node_fibers)))

    def update_prestress(self, u_nodes):
        """struct_update_prestress (PRESTRESS_TECH mulf) with the displacements of the finished step, (n_nodes, 3)"""
        lib.oracle_job_update_prestress(self.h, _p(_d(u_nodes)))

    def mulf_history(self):
        """(n_ele, ngp [+ 1], 18) inverse Jacobian | deformation gradient of every Gauss point [and the centroid: FORM_MULF_FBAR]"""
        lib.oracle_job_get_mulf_history.restype = C.c_longlong
        n = lib.oracle_job_get_mulf_history(self.h, None)
        out = np.zeros(n)
        lib.oracle_job_get_mulf_history(self.h, _p(out))
        return out.reshape(len(self.ele_gid), -1, 18)  # ngp records per element, MULF with F-bar: + the centroid

    def recover(self, du_nodes, step_length=1.0):
        """struct_calc_recover (EAS): alpha of every element += step_length * alpha_inc(Newton increment), (n_nodes, 3)"""
        lib.oracle_job_recover(self.h, _p(_d(du_nodes)), C.c_double(step_length))

    def eas_history(self):
        """(n_ele, 3 n + n^2 + 24 n + 1): alpha_inc | alpha | s | invKaa | Kda | old_step_length (empty unless FORM_EAS_*)"""
        lib.oracle_job_get_eas_history.restype = C.c_longlong
        n = lib.oracle_job_get_eas_history(self.h, None)
        out = np.zeros(n)
        lib.oracle_job_get_eas_history(self.h, _p(out))
        return out.reshape(len(self.ele_gid), -1) if n else out

    def set_eas_history(self, hist):
        lib.oracle_job_set_eas_history(self.h, _p(_d(hist)))

    def element_gp_stress(self, e, u_nodes, stress_type=1):
        """(ngp, 6) Gauss-point stresses of element e with the job's formulation, material, fibres and history;
        stress_type = OUTPUT_STRAIN + t: the strains of type t instead"""
        out = np.zeros((NGP[self.ct], 6))
        err = lib.oracle_job_element_gp_stress(self.h, C.c_int(e), _p(_d(u_nodes)), C.c_int(stress_type), _p(out))
        return err, out

    def rank(self, r):
        sizes = np.zeros(8, np.int64)
        lib.oracle_job_rank_sizes(self.h, C.c_int(r), _p(sizes))
        out = dict(n_row_nodes=int(sizes[0]), n_col_nodes=int(sizes[1]), n_row_eles=int(sizes[2]),
                   n_col_eles=int(sizes[3]), n_row_dof=int(sizes[4]), n_col_dof=int(sizes[5]), nnz=int(sizes[6]))
        out["col_node_gid"] = np.zeros(out["n_col_nodes"], np.int32)
        out["col_eles"] = np.zeros(out["n_col_eles"], np.int32)
        out["row_dof_gid"] = np.zeros(out["n_row_dof"], np.int32)
        out["col_dof_gid"] = np.zeros(out["n_col_dof"], np.int32)
        out["rowptr"] = np.zeros(out["n_row_dof"] + 1, np.int64)
        out["colind"] = np.zeros(out["nnz"], np.int32)
        lib.oracle_job_rank_maps(self.h, C.c_int(r), _p(out["col_node_gid"]), _p(out["col_eles"]),
                                 _p(out["row_dof_gid"]), _p(out["col_dof_gid"]), _p(out["rowptr"]), _p(out["colind"]))
        return out

    def set_dirichlet_rows(self, r, row_lids):
        row_lids = _i(row_lids)
        lib.oracle_job_set_dirichlet_rows(self.h, C.c_int(r), C.c_int(len(row_lids)), _p(row_lids))

    def set_u_nodal(self, u_nodes):
        """u_nodes: (n_nodes, 3) in the order of self.node_gid."""
        lib.oracle_job_set_u_nodal(self.h, _p(_d(u_nodes)))

    def set_u(self, r, u_row):
        lib.oracle_job_set_u(self.h, C.c_int(r), _p(_d(u_row)))

    def evaluate(self, action=ACTION_NLNSTIFF, max_threads=1):
        fb = C.c_int(-1)
        sec = C.c_double(0)
        bad = lib.oracle_job_evaluate(self.h, C.c_int(action), C.c_int(max_threads), C.byref(fb), C.byref(sec))
        self.seconds = sec.value
        return bad, fb.value

    def get_f(self, r, n):
        f = np.zeros(n)
        lib.oracle_job_get_f(self.h, C.c_int(r), _p(f))
        return f

    def get_K(self, r, nnz):
        K = np.zeros(nnz)
        lib.oracle_job_get_K(self.h, C.c_int(r), _p(K))
        return K

    def __del__(self):
        try:
            lib.oracle_job_destroy(self.h)
        except Exception:
            pass


def smooth_displacement(X, h, seed=12345, amp=0.02, noise=1e-3):
    """The synthetic displacement state of SURVEY §8d: affine + sine + uniform noise U(-noise,noise)*h."""
    rng = np.random.Generator(np.random.PCG64(seed))
    A = np.array([[0.03, 0.01, -0.02], [0.00, -0.02, 0.015], [0.01, 0.02, 0.025]])
    u = X @ A.T + amp * np.sin(2 * np.pi * X[:, [1, 2, 0]])
    u += rng.uniform(-noise, noise, size=X.shape) * h
    return u
