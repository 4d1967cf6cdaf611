"""Minimal stand-in for the caller of the path (src/structure_new: Newton loop, Dirichlet handling,
linear solve) used by the solve-level golden-vector tests.  scipy does the linear algebra; the
assembled K and f come either from the oracle (CPU tests) or from the CUDA library (GPU tests)."""
import numpy as np
import scipy.sparse as sp
import scipy.sparse.linalg as spla


def face_traction_forces(X, conn_faces, t):
    """Consistent nodal forces of a uniform traction t (force per reference area, 'Live' Neumann
    condition) on bilinear quad faces: 2x2 Gauss, N_a * t * dA."""
    f = np.zeros((X.shape[0], 3))
    g = 1.0 / np.sqrt(3.0)
    for quad in conn_faces:
        P = X[quad]
        for r in (-g, g):
            for s in (-g, g):
                N = 0.25 * np.array([(1 - r) * (1 - s), (1 + r) * (1 - s), (1 + r) * (1 + s), (1 - r) * (1 + s)])
                dr = 0.25 * np.array([-(1 - s), (1 - s), (1 + s), -(1 + s)])
                ds = 0.25 * np.array([-(1 - r), -(1 + r), (1 + r), (1 - r)])
                dA = np.linalg.norm(np.cross(dr @ P, ds @ P))
                f[quad] += np.outer(N, t) * dA
    return f


def hex8_faces_on_plane(X, conn, axis, value, tol=1e-12):
    """faces (4 node ids, hex8 local numbering) of elements lying in the plane X[axis] == value"""
    local = [(0, 3, 2, 1), (4, 5, 6, 7), (0, 1, 5, 4), (1, 2, 6, 5), (2, 3, 7, 6), (3, 0, 4, 7)]
    out = []
    for en in conn:
        for fnodes in local:
            ids = en[list(fnodes)]
            if np.all(np.abs(X[ids, axis] - value) < tol):
                out.append(ids)
    return np.array(out)


def solve_with_dirichlet(rowptr, colind, K, rhs, fixed_dofs):
    n = len(rowptr) - 1
    A = sp.csr_matrix((K, colind, rowptr), shape=(n, n)).tolil()
    rhs = rhs.copy()
    for d in fixed_dofs:
        A.rows[d], A.data[d] = [d], [1.0]
        rhs[d] = 0.0
    A = A.tocsc()
    # symmetric elimination is not needed for a direct solve
    return spla.spsolve(A, rhs)


def newton_statics(assemble, n_dof, f_ext, fixed_dofs, rowptr, colind, maxiter=20, tol=1e-10):
    """full Newton on f_int(u) - f_ext = 0.  assemble(u) -> (f_int, K_values).  Returns u, residual norms."""
    u = np.zeros(n_dof)
    norms = []
    for _ in range(maxiter):
        f_int, K = assemble(u)
        r = f_ext - f_int
        r[fixed_dofs] = 0.0
        norms.append(float(np.linalg.norm(r)))
        if norms[-1] < tol:
            break
        u += solve_with_dirichlet(rowptr, colind, K, r, fixed_dofs)
    return u, norms
