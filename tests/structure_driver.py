"""Mock of the caller of the path for the reference's solve-level golden vectors (tests/golden/*.json):
what src/structure_new does around Discretization::evaluate, restated in numpy for 1-rank problems.

  * surface Neumann loads, TYPE Live          src/solid_3D_ele/4C_solid_3D_ele_surface_evaluate.cpp:225-330
    (rules quad_4point / quad_9point / tri_6point with the reference's truncated constants,
     4C_fem_general_utils_integration.cpp:6151-6183, 6185-6231, 6374-6396; rule choice 4C_solid_3D_ele_surface.cpp:58-77)
  * statics: f_int(d_{n+1}) = f_ext(t_{n+1})  src/structure_new/src/implicit/4C_structure_new_impl_statics.cpp
  * GenAlpha, TR-like mid-averaging           .../implicit/4C_structure_new_impl_genalpha.cpp:255-296, 380-410, 468-485
    parameters from RHO_INF (default 1)       .../utils/4C_structure_new_utils.cpp:372-410
    consistent initial acceleration           .../4C_structure_new_integrator.cpp:175-300
    force assembly                            .../model_evaluator/4C_structure_new_model_evaluator_structure.cpp:244-262, 1486-1520
  * full Newton, Dirichlet rows replaced by identity (dense solve; the problems have <= 207 DOFs)

The element work (f_int, K, M) comes from a backend: the CPU oracle or the CUDA library.  Test infrastructure.
"""
import json
import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CT = {"HEX8": 0, "HEX27": 1, "TET10": 2, "HEX20": 3, "TET4": 4, "WEDGE6": 5, "PYRAMID5": 6, "HEX18": 7, "NURBS27": 8}
# faces of a nurbs27 element and the parent directions of their two knot vectors (4C_fem_general_utils_local_connectivity_matrices.hpp:250-253,
# 4C_fem_nurbs_discretization_knotvector.cpp:265-413)
NURBS27_FACES = [[0, 1, 2, 3, 4, 5, 6, 7, 8], [18, 19, 20, 21, 22, 23, 24, 25, 26], [0, 1, 2, 9, 10, 11, 18, 19, 20],
                 [6, 7, 8, 15, 16, 17, 24, 25, 26], [2, 5, 8, 11, 14, 17, 20, 23, 26], [0, 3, 6, 9, 12, 15, 18, 21, 24]]
NURBS27_FACE_DIRS = [(0, 1), (0, 1), (0, 2), (0, 2), (1, 2), (1, 2)]
# faces of the SURVEY 8f F4 cell types as (face type, local nodes in the node order of the surface cell); face types
# 0 quad4, 3 quad8, 4 tri3.  Node numbering of the faces: 4C_fem_general_utils_local_connectivity_matrices.hpp:51-86
FACE_TABLES = {
    3: [(3, f) for f in ([0, 3, 2, 1, 11, 10, 9, 8], [0, 1, 5, 4, 8, 13, 16, 12], [1, 2, 6, 5, 9, 14, 17, 13], [2, 3, 7, 6, 10, 15, 18, 14],
                         [0, 4, 7, 3, 12, 19, 15, 11], [4, 5, 6, 7, 16, 17, 18, 19])],
    4: [(4, f) for f in ([0, 1, 3], [1, 2, 3], [0, 3, 2], [0, 2, 1])],
    5: [(4, [2, 1, 0]), (4, [3, 4, 5]), (0, [0, 1, 4, 3]), (0, [1, 2, 5, 4]), (0, [2, 0, 3, 5])],
    6: [(0, [0, 3, 2, 1]), (4, [0, 1, 4]), (4, [1, 2, 4]), (4, [2, 3, 4]), (4, [0, 4, 3])],
    # hex18: two quad9 faces, four quad6 side faces (4C_fem_general_utils_local_connectivity_matrices.hpp:38-43)
    7: [(1, [0, 3, 2, 1, 7, 6, 5, 4, 8]), (1, [9, 10, 11, 12, 13, 14, 15, 16, 17]), (5, [0, 1, 4, 9, 10, 13]), (5, [1, 2, 5, 10, 11, 14]),
        (5, [2, 3, 6, 11, 12, 15]), (5, [3, 0, 7, 12, 9, 16])],
}


# the same for hex8 / hex27 / tet10 (:45-63), used where the orientation of a face matters (orthopressure loads: outward normals);
# their Live loads keep the orientation-free search of Problem._element_faces
ORIENTED_FACES = {
    0: [(0, f) for f in ([0, 3, 2, 1], [0, 1, 5, 4], [1, 2, 6, 5], [2, 3, 7, 6], [0, 4, 7, 3], [4, 5, 6, 7])],
    1: [(1, f) for f in ([0, 3, 2, 1, 11, 10, 9, 8, 20], [0, 1, 5, 4, 8, 13, 16, 12, 21], [1, 2, 6, 5, 9, 14, 17, 13, 22],
                         [2, 3, 7, 6, 10, 15, 18, 14, 23], [0, 4, 7, 3, 12, 19, 15, 11, 24], [4, 5, 6, 7, 16, 17, 18, 19, 25])],
    2: [(2, f) for f in ([0, 1, 3, 4, 8, 7], [1, 2, 3, 5, 9, 8], [0, 3, 2, 7, 9, 6], [0, 2, 1, 6, 5, 4])],
}


def load(name):
    with open(os.path.join(GOLDEN, name + ".json")) as f:
        return json.load(f)


class Problem:
    def __init__(self, fx, node_param_coords):
        """node_param_coords(ct) -> (nen, 3) parameter coordinates of the element nodes (reference numbering)"""
        self.fx = fx
        ids = sorted(int(k) for k in fx["nodes"])
        if not fx.get("part_of_mixed_mesh"):
            # nodes that no element uses carry no DOFs that anything couples to (inputs derived from a richer mesh keep them)
            used = {n for e in fx["elements"] for n in e["nodes"]}
            ids = [i for i in ids if i in used]
        self.ids = ids
        self.index = {nid: i for i, nid in enumerate(ids)}
        self.X = np.array([fx["nodes"][str(i)] for i in ids], float)
        self.node_gid = np.array([i - 1 for i in ids], np.int32)
        types = {e["type"] for e in fx["elements"]}
        assert len(types) == 1
        self.ct = CT[types.pop()]
        self.conn = np.array([[self.index[n] for n in e["nodes"]] for e in fx["elements"]], np.int32)
        self.ele_gid = np.array([e["id"] - 1 for e in fx["elements"]], np.int32)
        self.ele_knots = self.weights = None
        if self.ct == 8:
            # knot span of every element (Knotvector::get_ele_knots, 4C_fem_nurbs_discretization_knotvector.cpp:116-243): element
            # gids count through the patches, inside a patch u runs fastest; an element with cartesian ids (i, j, k) takes the
            # knots [i, i + 6) x [j, j + 6) x [k, k + 6) of its patch.  Control point weights: the fourth number of a CP line
            self.weights = np.array([fx["weights"][str(i)] for i in ids], float)
            kn, off = [], 0
            for patch in fx["knotvectors"]:
                assert all(k["degree"] == 2 and k["type"] == "Interpolated" for k in patch) and len(patch) == 3
                nele = [len(k["knots"]) - 5 for k in patch]
                for loc in range(nele[0] * nele[1] * nele[2]):
                    c = (loc % nele[0], (loc // nele[0]) % nele[1], loc // (nele[0] * nele[1]))
                    kn.append([patch[d]["knots"][c[d]:c[d] + 6] for d in range(3)])
                off += nele[0] * nele[1] * nele[2]
            kn = np.array(kn)
            self.ele_knots = kn[self.ele_gid - self.ele_gid.min()]
        m1 = fx["materials"]["1"]
        self.summands, self.fibers, self.node_fibers = None, None, None
        if m1[0] == "MAT_Struct_StVenantKirchhoff":
            m10, self.mat = m1, 0  # FC_MAT_STVENANTKIRCHHOFF
        else:
            assert m1[0] == "MAT_ElastHyper"
            nm = int(m1[m1.index("NUMMAT") + 1])
            mids = m1[m1.index("MATIDS") + 1:m1.index("MATIDS") + 1 + nm]
            m10 = fx["materials"][mids[0]]
            if nm == 1 and m10[0] == "ELAST_CoupNeoHooke":
                self.mat = 1  # FC_MAT_ELASTHYPER_COUPNEOHOOKE
            else:
                # summand list (kind, 4 parameters): 0 CoupNeoHooke / 1 CoupSVK (YOUNG, NUE), 2 CoupAnisoExpo (K1, K2, K1COMP, K2COMP)
                self.mat, self.summands = 2, []
                for mid in mids:
                    m = fx["materials"][mid]
                    g = lambda k: float(m[m.index(k) + 1])
                    if m[0] in ("ELAST_CoupNeoHooke", "ELAST_CoupSVK"):
                        self.summands.append((0 if m[0] == "ELAST_CoupNeoHooke" else 1, [g("YOUNG"), g("NUE"), 0.0, 0.0]))
                    else:
                        # INIT 1: FIBER1 of the element line; INIT 3: FIBER1 of the fibre nodes, interpolated to the Gauss points
                        assert m[0] == "ELAST_CoupAnisoExpo" and int(g("INIT")) in (1, 3) and int(g("FIBER_ID")) == 1
                        self.summands.append((2, [g("K1"), g("K2"), g("K1COMP"), g("K2COMP")]))
                        if int(g("INIT")) == 3:
                            self.node_fibers = np.array([fx["node_fibers"][str(i)] for i in ids], float)
                m10 = next(fx["materials"][mid] for mid in mids if "YOUNG" in fx["materials"][mid])
                if any(k == 2 for k, _ in self.summands) and self.node_fibers is None:
                    fib = []
                    for e in fx["elements"]:
                        t = e["opts"].split()
                        i = t.index("FIBER1")
                        fib.append([float(x) for x in t[i + 1:i + 4]])
                    self.fibers = np.array(fib)
        def tech(opts):
            if "PRESTRESS_TECH mulf" in opts:
                return 6 if "TECH fbar" in opts else 3
            if "KINEM linear" in opts:
                return 10 if "TECH eas_mild" in opts else (11 if "TECH eas_full" in opts else 2)
            return {"fbar": 1, "eas_mild": 4, "eas_full": 5, "shell_eas": 7, "shell_ans": 8, "shell_eas_ans": 9}.get(opts.split()[opts.split().index("TECH") + 1], 0) if "TECH" in opts.split() else 0
        techs = {tech(e["opts"]) for e in fx["elements"]}
        assert len(techs) == 1
        self.form = techs.pop()  # FC_FORM_DISPLACEMENT / _FBAR / _LINEAR / _MULF / _EAS_MILD / _EAS_FULL / _MULF_FBAR
        io = fx.get("io", {})
        self.stress_type = {"2pk": 0, "cauchy": 1}.get(io.get("STRUCT_STRESS", "cauchy").lower(), 1)
        self.strain_type = {"gl": 0, "ea": 1, "log": 2}.get(io.get("STRUCT_STRAIN", "none").lower())  # None: no strain output
        self.dens = float(m1[m1.index("DENS") + 1])
        self.young = float(m10[m10.index("YOUNG") + 1])
        self.nue = float(m10[m10.index("NUE") + 1])
        self.n_dof = 3 * len(ids)
        dyn = fx["dynamic"]
        self.dt, self.numstep = float(dyn["TIMESTEP"]), int(dyn["NUMSTEP"])
        self.dynamic = dyn.get("DYNAMICTYPE", "GenAlpha")  # the default of STRUCTURAL DYNAMIC (4C_inpar_structure.cpp:80-92)
        # inputs that stop the Newton iteration early (TOLRES / TOLDISP far above rounding) and still list results to 1e-12: the
        # result then depends on where the reference stopped, so the replay applies the reference's test (absolute L2 norms of
        # the residual and of the increment, combined with NORMCOMBI_RESFDISP) instead of iterating to rounding level
        self.ref_stop = None
        # (also the "linear" tests that cut Newton off: MAXITER <= 3 with DIVERCONT continue, whatever the tolerances)
        cut_off = dyn.get("DIVERCONT", "stop") == "continue" and int(dyn.get("MAXITER", 50)) <= 3
        if (float(dyn.get("TOLRES", 0)) >= 1e-9 or cut_off) and dyn.get("NORM_RESF", "Abs") == "Abs" and dyn.get("NORM_DISP", "Abs") == "Abs":
            self.ref_stop = (float(dyn["TOLRES"]), float(dyn["TOLDISP"]), dyn.get("NORMCOMBI_RESFDISP", "And"))
        # MAXITER with DIVERCONT continue: the reference goes on with what it has after MAXITER iterations
        self.ref_maxiter = int(dyn["MAXITER"]) if self.ref_stop is not None and dyn.get("DIVERCONT", "stop") == "continue" and "MAXITER" in dyn else None
        # PRESTRESS mulf: steps with t <= PRESTRESSTIME end with struct_update_prestress and a reset of the displacements
        self.prestress_time = float(dyn["PRESTRESSTIME"]) if dyn.get("PRESTRESS", "none").lower() == "mulf" else None
        # Dirichlet: zero in most inputs; prescribed displacements VAL * FUNCT(X, t) are imposed at the start of every load
        # step (statics only)
        # Hierarchy of Core::FE::Utils::Dbc::read_dirichlet_condition (4C_fem_discretization_utils_dbc.cpp:164-400): surface, then
        # line, then point conditions; a condition sets its ONOFF = 1 DOFs, and its ONOFF = 0 releases a DOF only if a condition of
        # higher dimension fixed it (conditions of the same kind add up)
        toggle, level = np.zeros(self.n_dof, bool), np.full(self.n_dof, 9)
        prescribed = {}  # DOF -> (VAL, FUNCT id) of the condition that fixed it last
        for order, key, topo in ((2, "dirich", "dsurf"), (1, "dirich_line", "dline"), (0, "dirich_point", "dnode")):
            for c in fx.get(key, []):
                assert not any(c["val"]) or dyn.get("DYNAMICTYPE", "GenAlpha") == "Statics"
                for n in (m for m in fx[topo][str(c["E"])] if m in self.index):
                    for d in range(3):
                        lid = 3 * self.index[n] + d
                        if c["onoff"][d]:
                            toggle[lid], level[lid] = True, order
                            prescribed[lid] = (c["val"][d], c["funct"][d])
                        elif order < level[lid]:
                            toggle[lid], level[lid] = False, order
        self.fixed = np.nonzero(toggle)[0].astype(np.int64)
        self.prescribed = [(lid, v, f) for lid, (v, f) in sorted(prescribed.items()) if toggle[lid] and v != 0.0]
        self.xi_nodes = np.asarray(node_param_coords(self.ct), float)
        self.faces = []
        self.has_ortho = any(c["type"] != "Live" for c in fx["neumann"])
        self.has_follower = any(c["type"] == "orthopressure" for c in fx["neumann"])  # loads that depend on the current iterate
        self.loadlin = fx["dynamic"].get("LOADLIN", "no").lower() in ("yes", "true", "1")
        face_tables = {**FACE_TABLES, **ORIENTED_FACES} if self.has_ortho else FACE_TABLES
        for c in fx["neumann"]:
            # Live, or pseudo_orthopressure: pressure VAL[0] along the normal of the last converged configuration
            # or orthopressure: along the normal of the current iterate, with its load stiffness when LOADLIN is yes
            assert c["type"] in ("Live", "pseudo_orthopressure", "orthopressure") and not any(c["onoff"][3:])
            sset = {self.index[n] for n in fx["dsurf"][str(c["E"])] if n in self.index}
            for e, en in enumerate(self.conn):
                if self.ct == 8:
                    for sid, loc in enumerate(NURBS27_FACES):
                        if all(int(en[a]) in sset for a in loc):
                            self.faces.append((c, (e, sid), 6))  # nurbs9 face: named by parent element and surface number
                    continue
                if self.ct in face_tables:
                    for ft, loc in face_tables[self.ct]:
                        if all(int(en[a]) in sset for a in loc):
                            self.faces.append((c, en[loc], ft))  # already in the surface cell's node order
                    continue
                for loc, rs in self._element_faces():
                    if all(int(en[a]) in sset for a in loc):
                        self.faces.append((c, en[loc], rs))
        assert self.faces or not fx["neumann"] or fx.get("part_of_mixed_mesh")  # (the loaded faces may all belong to the other cell type)
        # the same loads as lists for a Neumann backend (oracle / CUDA library): faces by their nodes in the reference's
        # quad4 / quad9 / tri6 node order, volume conditions by their elements
        self.neumann = None  # backend with surface(), volume(), *_coords(); None = the numpy restatement below (surfaces only)
        self.face_lists, self.vol_lists = [], []
        for c in fx["neumann"]:
            if self.ct in face_tables or self.ct == 8:
                for ft in sorted({ft for cc, _, ft in self.faces if cc is c}):
                    self.face_lists.append((c, ft, np.array([nodes for cc, nodes, f in self.faces if cc is c and f == ft], np.int32)))
                continue
            fl = [self._ordered_face(nodes, rs) for cc, nodes, rs in self.faces if cc is c]
            if not fl:  # part of a mixed mesh: the faces of this condition belong to another group
                continue
            self.face_lists.append((c, {4: 0, 9: 1, 6: 2}[len(fl[0])], np.array(fl, np.int32)))
        for c in fx.get("vol_neumann", []):
            assert c["type"] == "Live" and not any(c["onoff"][3:])
            vset = {self.index[n] for n in fx["dvol"][str(c["E"])] if n in self.index}
            els = [e for e, en in enumerate(self.conn) if all(int(a) in vset for a in en)]
            assert els
            self.vol_lists.append((c, np.array(els, np.int32)))
        self._gp_cache = {}

    # faces of the parent cell as (local node ids, in-face parameter coordinates of those nodes)
    def _element_faces(self):
        xi = self.xi_nodes
        out = []
        if self.ct in (0, 1):
            for ax in range(3):
                for side in (-1.0, 1.0):
                    loc = [a for a in range(len(xi)) if xi[a, ax] == side]
                    other = [d for d in range(3) if d != ax]
                    out.append((loc, xi[loc][:, other]))
        else:
            corners = xi[:4]
            for skip in range(4):
                tri = [c for c in range(4) if c != skip]
                # barycentric coordinates of every node w.r.t. the three corners of this face
                loc, lam = [], []
                for a in range(10):
                    A = np.vstack([corners[tri].T, np.ones(3)])
                    l, res, *_ = np.linalg.lstsq(A, np.append(xi[a], 1.0), rcond=None)
                    if np.allclose(A @ l, np.append(xi[a], 1.0), atol=1e-12) and np.all(l > -1e-12):
                        loc.append(a)
                        lam.append(l)
                assert len(loc) == 6
                out.append((loc, np.array(lam)))
        return out

    def _ordered_face(self, nodes, rs):
        """face nodes in the node order of the reference's surface cell (4C_fem_general_utils_fem_shapefunctions.hpp:2003-2111)"""
        rs = np.asarray(rs, float)
        if self.ct in (0, 1):
            want = [(-1, -1), (1, -1), (1, 1), (-1, 1), (0, -1), (1, 0), (0, 1), (-1, 0), (0, 0)][:len(nodes)]
        else:
            want = [(1, 0, 0), (0, 1, 0), (0, 0, 1), (.5, .5, 0), (0, .5, .5), (.5, 0, .5)]
        out = []
        for w in want:
            k = [i for i in range(len(nodes)) if np.allclose(rs[i], w, atol=1e-12)]
            assert len(k) == 1
            out.append(int(nodes[k[0]]))
        return out

    def _surface_shape(self, rs_nodes, pt):
        """shape functions + derivatives w.r.t. the two surface parameters at pt, for the nodes of a face"""
        if self.ct in (0, 1):
            def lag(a, x):
                if self.ct == 0:
                    return 0.5 * (1 + a * x), 0.5 * a
                if a == 0:
                    return 1 - x * x, -2 * x
                return 0.5 * x * (x + a), x + 0.5 * a
            N, dr, ds = [], [], []
            for ra, sa in rs_nodes:
                Lr, dLr = lag(ra, pt[0])
                Ls, dLs = lag(sa, pt[1])
                N.append(Lr * Ls), dr.append(dLr * Ls), ds.append(Lr * dLs)
            return np.array(N), np.array(dr), np.array(ds)
        # tri6: rs_nodes holds barycentric coordinates (l0,l1,l2) of the nodes; pt = (r,s), l = (1-r-s, r, s)
        l = np.array([1 - pt[0] - pt[1], pt[0], pt[1]])
        dl = np.array([[-1.0, -1.0], [1.0, 0.0], [0.0, 1.0]])
        N, dr, ds = [], [], []
        for lam in rs_nodes:
            nz = [i for i in range(3) if lam[i] > 1e-9]
            if len(nz) == 1:
                i = nz[0]
                N.append(l[i] * (2 * l[i] - 1))
                d = (4 * l[i] - 1) * dl[i]
            else:
                i, j = nz
                N.append(4 * l[i] * l[j])
                d = 4 * (l[i] * dl[j] + l[j] * dl[i])
            dr.append(d[0]), ds.append(d[1])
        return np.array(N), np.array(dr), np.array(ds)

    def _surface_rule(self):
        if self.ct == 0:
            g = 0.5773502691896
            return [((-g, -g), 1.0), ((g, -g), 1.0), ((g, g), 1.0), ((-g, g), 1.0)]
        if self.ct == 1:
            g, w1, w2 = 0.7745966692415, 0.5555555555556, 0.8888888888889
            pts = [(-g, -g), (g, -g), (g, g), (-g, g), (0.0, -g), (g, 0.0), (0.0, g), (-g, 0.0), (0.0, 0.0)]
            return [(p, (w2 if p[0] == 0 else w1) * (w2 if p[1] == 0 else w1)) for p in pts]
        a, b, c, d = 0.0915762135098, 0.8168475729805, 0.4459484909160, 0.1081030181681
        wa, wc = 0.0549758718277, 0.1116907948390
        return [((a, a), wa), ((b, a), wa), ((a, b), wa), ((c, d), wc), ((c, c), wc), ((d, c), wc)]

    def funct(self, fid, x, t):
        if fid <= 0:
            return 1.0
        expr = self.fx["functs"][str(fid)]
        env = {"x": x[0], "y": x[1], "z": x[2], "t": t}
        for name, v in self.fx.get("funct_vars", {}).get(str(fid), {}).items():  # VARIABLE ... TYPE linearinterpolation
            env[name] = float(np.interp(t, v["times"], v["values"]))
        return float(eval(expr, {"__builtins__": {}}, env))

    def _factors(self, key, cond, coords, t):
        """FUNCT of the condition at the Gauss points: (n, ngp, 3), None when every factor is 1"""
        if not any(cond["onoff"][d] and cond["funct"][d] > 0 for d in range(3)):
            return None
        fac = np.ones(coords.shape)
        for d in range(3):
            if cond["onoff"][d] and cond["funct"][d] > 0:
                fac[..., d] = [[self.funct(cond["funct"][d], x, t) for x in row] for row in coords]
        return fac

    def f_ext(self, t, d_conv=None):
        """d_conv: the last converged displacements (the configuration of pseudo_orthopressure loads)"""
        if self.neumann is not None:
            # Neumann backend (oracle restatement or the CUDA library): the driver only evaluates the FUNCT expressions at the
            # Gauss-point coordinates the backend hands out
            nb = self.neumann
            f = np.zeros(self.n_dof)
            for i, (c, ft, fn) in enumerate(self.face_lists):
                xg = self._gp_cache.setdefault(("s", i), nb.surface_coords(ft, fn))
                if c["type"] in ("pseudo_orthopressure", "orthopressure"):  # d_conv: last converged state / current iterate
                    nb.surface_orthopressure(ft, fn, d_conv, c["val"][0], self._factors(("s", i), c, xg, t), f)
                    continue
                nb.surface(ft, fn, c["onoff"], c["val"], self._factors(("s", i), c, xg, t), f)
            for i, (c, els) in enumerate(self.vol_lists):
                xg = self._gp_cache.setdefault(("v", i), nb.volume_coords(els))
                nb.volume(els, c["onoff"], c["val"], self._factors(("v", i), c, xg, t), f)
            return f
        assert not self.vol_lists and self.ct not in FACE_TABLES, "the numpy restatement covers surface loads of hex8 / hex27 / tet10 only"
        f = np.zeros((len(self.ids), 3))
        for c, nodes, rs in self.faces:
            P = self.X[nodes]
            for pt, w in self._surface_rule():
                N, dr, ds = self._surface_shape(rs, pt)
                g1, g2 = dr @ P, ds @ P
                m11, m22, m12 = g1 @ g1, g2 @ g2, g1 @ g2
                detA = np.sqrt(m11 * m22 - m12 * m12)
                xgp = N @ P
                for dof in range(3):
                    if c["onoff"][dof]:
                        fac = w * detA * c["val"][dof] * self.funct(c["funct"][dof], xgp, t)
                        f[nodes, dof] += N * fac
        return f.reshape(-1)

    def k_load(self, t, d):
        """load stiffness of the orthopressure loads at the iterate d (LOADLIN yes), dense"""
        K = np.zeros((self.n_dof, self.n_dof))
        for i, (c, ft, fn) in enumerate(self.face_lists):
            if c["type"] == "orthopressure":
                xg = self._gp_cache.setdefault(("s", i), self.neumann.surface_coords(ft, fn))
                K += self.neumann.surface_orthopressure_linearization(ft, fn, d, c["val"][0], self._factors(("s", i), c, xg, t))
        return K

    def result_error(self, d):
        """max (|d - VALUE| / TOLERANCE) over the displacement results, and the number compared"""
        err, n = 0.0, 0
        comp = {"dispx": 0, "dispy": 1, "dispz": 2}
        for r in self.fx["results"]:
            if r["quantity"] in comp:
                v = d[3 * self.index[r["node"]] + comp[r["quantity"]]]
                err = max(err, abs(v - r["value"]) / r["tolerance"])
                n += 1
        return err, n


class OracleNeumann:
    """Neumann backend on the CPU oracle (oracle/fc_oracle_neumann.hpp)"""

    def __init__(self, oracle, p):
        self.o, self.p = oracle, p

    def _nurbs_faces(self, fn, onoff, val, fac, want_coords):
        """faces (parent element, surface number) of a nurbs27 mesh, one oracle call per face with its knots and weights"""
        p, out = self.p, []
        for i, (e, sid) in enumerate(fn):
            nodes = p.conn[e][NURBS27_FACES[sid]]
            self.o.set_nurbs_element(p.ele_knots[e][list(NURBS27_FACE_DIRS[sid])], p.weights[nodes])
            r = self.o.surface_neumann(6, p.X[nodes][None], onoff, val, None if fac is None else fac[i][None], want_coords=want_coords)
            out.append((nodes, r[1][0] if want_coords else r[0]))
        return out

    def surface_coords(self, ft, fn):
        if ft == 6:
            return np.array([xg for _, xg in self._nurbs_faces(fn, [0, 0, 0], [0.0, 0.0, 0.0], None, True)])
        return self.o.surface_neumann(ft, self.p.X[fn], [0, 0, 0], [0.0, 0.0, 0.0], want_coords=True)[1]

    def volume_coords(self, els):
        return self.o.volume_neumann(self.p.ct, self.p.X[self.p.conn[els]], [0, 0, 0], [0.0, 0.0, 0.0], want_coords=True)[1]

    def surface_orthopressure(self, ft, fn, d, value, fac, f):
        xc = self.p.X[fn] + (0.0 if d is None else d.reshape(-1, 3)[fn])
        np.add.at(f.reshape(-1, 3), fn, self.o.surface_neumann_orthopressure(ft, xc, value, fac))

    def surface_orthopressure_linearization(self, ft, fn, d, value, fac):
        n = self.p.n_dof
        K = np.zeros((n, n))
        Ke = self.o.surface_neumann_orthopressure_linearization(ft, self.p.X[fn] + d.reshape(-1, 3)[fn], value, fac)
        for nodes, ke in zip(fn, Ke):
            dofs = (3 * nodes[:, None] + np.arange(3)).reshape(-1)
            K[np.ix_(dofs, dofs)] += ke
        return K

    def surface(self, ft, fn, onoff, val, fac, f):
        if ft == 6:
            for nodes, fe in self._nurbs_faces(fn, onoff, val, fac, False):
                np.add.at(f.reshape(-1, 3), nodes, fe)
            return
        np.add.at(f.reshape(-1, 3), fn, self.o.surface_neumann(ft, self.p.X[fn], onoff, val, fac))

    def volume(self, els, onoff, val, fac, f):
        np.add.at(f.reshape(-1, 3), self.p.conn[els], self.o.volume_neumann(self.p.ct, self.p.X[self.p.conn[els]], onoff, val, fac))


class CudaNeumann:
    """Neumann backend on the CUDA library (fc_solid_neumann_*): f_ext is assembled on the device"""

    def __init__(self, solid):
        self.s = solid

    def surface_coords(self, ft, fn):
        if ft == 6:
            return self.s.neumann_surface_nurbs_gp_coords(fn[:, 0], fn[:, 1])
        return self.s.neumann_surface_gp_coords(ft, fn)

    def volume_coords(self, els):
        return self.s.neumann_volume_gp_coords(els)

    def surface_orthopressure(self, ft, fn, d, value, fac, f):
        self.s.neumann_surface_orthopressure(ft, fn, d, value, f, gp_factor=fac)

    def surface_orthopressure_linearization(self, ft, fn, d, value, fac):
        rowptr, colind = self.s.get_graph()
        K = np.zeros(len(colind))
        self.s.neumann_surface_orthopressure_linearization(ft, fn, d, value, K, gp_factor=fac)
        return csr_to_dense(rowptr, colind, K, len(rowptr) - 1)

    def surface(self, ft, fn, onoff, val, fac, f):
        if ft == 6:
            return self.s.neumann_surface_nurbs(fn[:, 0], fn[:, 1], onoff, val, f, gp_factor=fac)
        self.s.neumann_surface(ft, fn, onoff, val, f, gp_factor=fac)

    def volume(self, els, onoff, val, fac, f):
        self.s.neumann_volume(els, onoff, val, f, gp_factor=fac)


def _solve(A, rhs, fixed):
    A = A.copy()
    rhs = rhs.copy()
    A[fixed, :] = 0.0
    A[fixed, fixed] = 1.0
    rhs[fixed] = 0.0
    if len(rhs) > 1500:  # the larger meshes: sparse LU instead of a dense one (minutes -> seconds)
        import scipy.sparse as sp
        import scipy.sparse.linalg as spla
        return spla.splu(sp.csc_matrix(A), permc_spec="MMD_AT_PLUS_A", diag_pivot_thresh=0.0, options=dict(SymmetricMode=True)).solve(rhs)
    return np.linalg.solve(A, rhs)


def _converged(norms, inc, newton_tol):
    """to rounding level: tiny residual and increment, or a residual that sits on its rounding floor (dropped by
    1e-9 from its peak and stopped decreasing) - the reference stops much earlier (TOLRES/TOLDISP)"""
    if norms[-1] < 1e-10 and np.linalg.norm(inc) < newton_tol:
        return True
    return len(norms) > 2 and norms[-1] < 1e-9 * max(norms) and norms[-1] > 0.5 * norms[-2]


def run(problem, assemble, mass=None, newton_tol=1e-13, maxiter=30, solve=None, update_prestress=None, recover=None):
    """assemble(d) -> (f_int [n], K dense [n, n]); mass() -> M dense.  Returns d at MAXTIME and the residual
    norm history of every step (list of lists).  solve(A, rhs, fixed) replaces the dense direct solve (A is then
    whatever assemble/mass return, e.g. device-resident CSR values supporting a*A + b*B and A @ x).
    update_prestress(d): the element action of a MULF prestress step (problems with PRESTRESS mulf): after a step with
    t <= PRESTRESSTIME the elements update their history and the displacements are reset
    (Solid::ModelEvaluator::Structure::update_step_element, 4C_structure_new_model_evaluator_structure.cpp:1555-1592;
    Solid::IMPLICIT::PreStress::post_update, 4C_structure_new_impl_prestress.cpp).
    recover(inc): struct_calc_recover after every Newton update (elements with condensed variables, EAS:
    Solid::ModelEvaluator::Structure::run_recover, 4C_structure_new_model_evaluator_structure.cpp)."""
    _solve = solve or globals()["_solve"]
    p = problem
    n, fixed, dt = p.n_dof, p.fixed, p.dt
    d = np.zeros(n)
    history = []
    if p.dynamic == "Statics":
        for step in range(1, p.numstep + 1):
            fext = p.f_ext(step * dt, d.copy()) if p.has_ortho else p.f_ext(step * dt)
            norms = []
            for lid, v, fid in p.prescribed:  # prescribed displacements of this step
                d[lid] = v * p.funct(fid, p.X[lid // 3], step * dt)
            if p.ref_stop is not None:
                # the reference's loop: solve, update, recover, new residual, test (converged if the residual OR / AND the increment
                # is below its tolerance)
                tolres, toldisp, combi = p.ref_stop

                def system(dd):  # follower loads: external force and load stiffness of the iterate
                    fi, Kt = assemble(dd)
                    fe = p.f_ext(step * dt, dd) if p.has_follower else fext
                    if p.has_follower and p.loadlin:
                        Kt = Kt + p.k_load(step * dt, dd)
                    return fi, Kt, fe
                fint, K, fe = system(d)
                r = fint - fe
                r[fixed] = 0.0
                for _ in range(p.ref_maxiter or maxiter):
                    norms.append(float(np.linalg.norm(r)))
                    inc = _solve(K, -r, fixed)
                    d = d + inc
                    if recover is not None:
                        recover(inc)
                    fint, K, fe = system(d)
                    r = fint - fe
                    r[fixed] = 0.0
                    a, b = np.linalg.norm(r) < tolres, np.linalg.norm(inc) < toldisp
                    if (a or b) if combi == "Or" else (a and b):
                        break
                history.append(norms)
                if p.prestress_time is not None and step * dt <= p.prestress_time + 1.0e-15:
                    update_prestress(d)
                    d = np.zeros(n)
                continue
            for _ in range(maxiter):
                fint, K = assemble(d)
                if p.has_follower:
                    fext = p.f_ext(step * dt, d)
                    if p.loadlin:
                        K = K + p.k_load(step * dt, d)
                r = fint - fext
                r[fixed] = 0.0
                norms.append(float(np.linalg.norm(r)))
                inc = _solve(K, -r, fixed)
                d = d + inc
                if recover is not None:
                    recover(inc)
                if _converged(norms, inc, newton_tol):
                    break
            history.append(norms)
            if p.prestress_time is not None and step * dt <= p.prestress_time + 1.0e-15:
                update_prestress(d)
                d = np.zeros(n)
        return d, history

    assert p.dynamic == "GenAlpha"
    rho_inf = 1.0
    am = (2.0 * rho_inf - 1.0) / (rho_inf + 1.0)
    af = rho_inf / (rho_inf + 1.0)
    beta = 0.25 * (1.0 - am + af) ** 2
    gamma = 0.5 - am + af
    M = mass()
    v, t = np.zeros(n), 0.0
    fint_n, _ = assemble(d)
    fext_n = p.f_ext(t)
    a = _solve(M, -(fint_n - fext_n), fixed)  # consistent initial acceleration, zero on Dirichlet DOFs
    for step in range(1, p.numstep + 1):
        t_np = step * dt
        fext_np = p.f_ext(t_np)
        fstruct_old = af * (fint_n - fext_n)
        finert_n = M @ a
        cv = (beta - gamma) / beta * v + (2.0 * beta - gamma) * dt / (2.0 * beta) * a - gamma / (beta * dt) * d
        ca = (2.0 * beta - 1.0) / (2.0 * beta) * a - 1.0 / (beta * dt) * v - 1.0 / (beta * dt * dt) * d
        d_np = d.copy()  # PREDICT ConstDis
        norms = []
        for _ in range(maxiter):
            a_np = ca + d_np / (beta * dt * dt)
            fint, K = assemble(d_np)
            r = (1.0 - af) * (fint - fext_np) + fstruct_old + (1.0 - am) * (M @ a_np) + am * finert_n
            r[fixed] = 0.0
            norms.append(float(np.linalg.norm(r)))
            J = (1.0 - af) * K + (1.0 - am) / (beta * dt * dt) * M
            inc = _solve(J, -r, fixed)
            d_np = d_np + inc
            if recover is not None:
                recover(inc)
            if _converged(norms, inc, newton_tol):
                break
        history.append(norms)
        a_np = ca + d_np / (beta * dt * dt)
        v_np = cv + gamma / (beta * dt) * d_np
        fint_n, _ = assemble(d_np)
        fext_n = fext_np
        d, v, a, t = d_np, v_np, a_np, t_np
    return d, history


def csr_to_dense(rowptr, colind, vals, n):
    A = np.zeros((n, n))
    for r in range(n):
        s, e = rowptr[r], rowptr[r + 1]
        A[r, colind[s:e]] = vals[s:e]
    return A


def extrapolation_matrix(ct, gauss_xi, shape, xi_nodes):
    """Gauss points -> nodes (4C_fem_general_utils_gauss_point_extrapolation.cpp:29-44, 90-260, 288-320): invert
    M = [N^base_i(xi_gp)] of the base cell (hex8/8 gp -> hex8, hex27/27 gp -> hex27, tet10/4 gp -> tet4), then interpolate the
    remaining nodes with the base cell's shape functions.  shape(ct, xi) -> N of the full cell."""
    if ct in (0, 1, 5):  # hex8 / hex27 / wedge6: the cell is its own base cell
        M = np.array([shape(ct, x) for x in gauss_xi])  # [gp][node]
        return np.linalg.inv(M)  # [node][gp]
    if ct == 3:  # hex20 with 27 Gauss points: base cell hex27, the 20 nodes are its first 20 (:29-44, :211-262)
        M = np.array([shape(1, x) for x in gauss_xi])
        return np.linalg.inv(M)[:20]
    if ct == 7:
        # hex18 with 18 Gauss points: base cell hex8 by least squares (M^T M)^-1 M^T (:150-208), then the hex8 functions at the nodes
        M = np.array([shape(0, x) for x in gauss_xi])  # [18][8]
        P = np.linalg.inv(M.T @ M) @ M.T
        return np.array([shape(0, x) for x in xi_nodes]) @ P
    if ct == 4:  # tet4 with one Gauss point: base "cell" point1, every node takes the Gauss-point value
        return np.ones((4, 1))
    if ct == 6:
        # pyramid5: the reference's hard-coded 5 x 8 matrix (:316-378: "the 8 Gauss points create a HEX-element inside the
        # pyramid"), used with the Gauss-point order of pyramid_8point as it stands
        a, b, c, d, e, f, g, h = (2.408235313815748, -0.6452847075210328, 0.1729035162684118, -0.542209910031327, 0.1452847075210439,
                                  -0.03892892005285509, -0.2702847075210531, 0.520284707521053)
        lo, hi = [a, b, c, b], [d, e, f, e]
        E = np.zeros((5, 8))
        for i in range(4):
            for k in range(4):
                E[i, k] = lo[(k - i) % 4]
                E[i, 4 + k] = hi[(k - i) % 4]
        E[4, :4], E[4, 4:] = g, h
        return E
    tet4 = lambda x: np.array([1.0 - x[0] - x[1] - x[2], x[0], x[1], x[2]])  # noqa: E731
    M = np.array([tet4(x) for x in gauss_xi])
    base = np.linalg.inv(M)  # [4][gp]
    B = np.array([tet4(x) for x in xi_nodes])  # [10][4]
    return B @ base


def nodal_stresses(problem, gp_stress_of_element, E):
    """assemble_extrapolated_nodal_values with nodal_average = true: every adjacent element adds
    (E gp_stress)[a] / (number of elements at the node).  gp_stress_of_element(e) -> (ngp, 6)."""
    p = problem
    nn = len(p.ids)
    count = np.zeros(nn)
    for en in p.conn:
        count[en] += 1
    out = np.zeros((nn, 6))
    for e, en in enumerate(p.conn):
        out[en] += (E @ gp_stress_of_element(e)) / count[en][:, None]
    return out


STRESS_COMP = {"stress_xx": 0, "stress_yy": 1, "stress_zz": 2, "stress_xy": 3, "stress_yz": 4, "stress_xz": 5}


def stress_result_error(problem, nodal, prefix="stress"):
    """max (|nodal stress - VALUE| / TOLERANCE) over the stress (prefix "strain": strain) results of the fixture, and the
    number compared"""
    err, n = 0.0, 0
    comp = {k.replace("stress", prefix): v for k, v in STRESS_COMP.items()}
    for r in problem.fx["results"]:
        if r["quantity"] in comp:
            v = nodal[problem.index[r["node"]], comp[r["quantity"]]]
            err = max(err, abs(v - r["value"]) / r["tolerance"])
            n += 1
    return err, n
