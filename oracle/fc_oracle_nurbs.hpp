// ORACLE — TEST INFRASTRUCTURE ONLY. Never linked into or called by the product library.
//
// NURBS basis of the nurbs27 solid element and its nurbs9 faces, restated from (paths relative to /root/reference)
//   B-spline value + derivative   src/core/fem/src/general/nurbs/4C_fem_general_utils_bspline.hpp:539-760
//                                 (Cox - de Boor on the element's local knot vector of 2 p + 2 knots, factors over
//                                 knot spans below 10e-9 set to zero)
//   3D functions / derivatives    src/core/fem/src/general/nurbs/4C_fem_general_utils_nurbs_shapefunctions.hpp:2586-2931
//   2D (faces)                    the same file, nurbs_get_2d_funct_deriv
//   element knots                 src/core/fem/src/nurbs_discretization/4C_fem_nurbs_discretization_knotvector.cpp:173-243
//   face knots, normal factor     the same file :251-420
// The reference hands knots and control point weights to the shape functions through ElementNodes
// (src/solid_3D_ele/4C_solid_3D_ele_calc_lib.hpp:112-150, 193-200, 397-404); the oracle keeps the current element's
// 3 x 6 knots and 27 weights in a thread-local context that the element loop sets before it evaluates the element.
#pragma once
#include <cmath>

namespace fc_oracle
{
  struct NurbsContext
  {
    const double* knots = nullptr;    // [3][6] of the current element (or [2][6] of a face)
    const double* weights = nullptr;  // [27] (or [9])
  };
  inline NurbsContext& nurbs_context()
  {
    static thread_local NurbsContext c;
    return c;
  }

  // value and derivative of local basis function ldofid (0..degree) of degree p at x          bspline.hpp:539-760
  inline void bspline_and_deriv(const double* kn, int p, double x, int ldofid, double& val, double& der)
  {
    double b[4] = {0.0, 0.0, 0.0, 0.0};
    b[p - ldofid] = 1.0;
    auto ratio = [](double num, double den) { return std::fabs(den) < 10e-9 ? 0.0 : num / den; };
    for (int q = 0; q < p - 1; ++q)
      for (int rr = 0; rr < p - q; ++rr)
      {
        const int i = ldofid + rr;
        const double f1 = ratio(x - kn[i], kn[i + q + 1] - kn[i]);
        const double f2 = ratio(kn[i + q + 2] - x, kn[i + q + 2] - kn[i + 1]);
        b[rr] = f1 * b[rr] + f2 * b[rr + 1];
      }
    const double d1 = kn[ldofid + p] - kn[ldofid], d2 = kn[ldofid + p + 1] - kn[ldofid + 1];
    val = ratio(x - kn[ldofid], d1) * b[0] + ratio(kn[ldofid + p + 1] - x, d2) * b[1];
    der = ratio((double)p, d1) * b[0] - ratio((double)p, d2) * b[1];
  }

  // N [n], dN [dim][n] column-major (dN[d + dim * a]) at uv in [-1, 1]^dim; n = 3^dim control points, u fastest
  template <int DIM>
  inline bool nurbs_funct_deriv(const double* uv, const double* knots, const double* weights, double* N, double* dN)
  {
    constexpr int P = 2, NB = DIM == 3 ? 27 : 9;
    double Jd[3] = {1, 1, 1}, val[3][3], der[3][3];
    for (int d = 0; d < DIM; ++d)
    {
      const double* kn = knots + 6 * d;
      Jd[d] = (kn[P + 1] - kn[P]) / 2.0;
      const double x = kn[P] + (uv[d] + 1.0) * Jd[d];
      for (int k = 0; k <= P; ++k) bspline_and_deriv(kn, P, x, k, val[d][k], der[d][k]);
    }
    double B[NB], dB[3][NB], sw = 0.0, sdw[3] = {0, 0, 0};
    for (int id = 0; id < NB; ++id)
    {
      const int rr = id % 3, mm = (id / 3) % 3, nn = id / 9;
      const double vn = DIM == 3 ? val[2][nn] : 1.0, dn = DIM == 3 ? der[2][nn] : 0.0;
      B[id] = val[0][rr] * val[1][mm] * vn;
      dB[0][id] = der[0][rr] * val[1][mm] * vn;
      dB[1][id] = val[0][rr] * der[1][mm] * vn;
      dB[2][id] = val[0][rr] * val[1][mm] * dn;
    }
    for (int id = 0; id < NB; ++id)
    {
      sw += weights[id] * B[id];
      for (int d = 0; d < DIM; ++d) sdw[d] += weights[id] * dB[d][id];
    }
    if (sw == 0.0) return false;
    for (int id = 0; id < NB; ++id)
    {
      if (N) N[id] = weights[id] * B[id] / sw;
      if (dN)
        for (int d = 0; d < DIM; ++d)
        {
          double v = dB[d][id] * sw - B[id] * sdw[d];
          v *= weights[id] / (sw * sw);
          dN[d + DIM * id] = v * Jd[d];
        }
    }
    return true;
  }

  // parameter coordinates of the 27 control points' positions in the element cube            local_connectivity_matrices.hpp:305-314
  inline void nurbs27_node_param(int a, double* xi)
  {
    xi[0] = (double)(a % 3 - 1);
    xi[1] = (double)((a / 3) % 3 - 1);
    xi[2] = (double)(a / 9 - 1);
  }

  // faces of a nurbs27 element (eleNodeNumbering_nurbs27_surfaces, local_connectivity_matrices.hpp:250-253) and the parent
  // directions their two knot vectors come from (knotvector.cpp:265-413)
  inline const int* nurbs27_face_nodes(int face)
  {
    static const int f[6][9] = {{0, 1, 2, 3, 4, 5, 6, 7, 8}, {18, 19, 20, 21, 22, 23, 24, 25, 26}, {0, 1, 2, 9, 10, 11, 18, 19, 20},
        {6, 7, 8, 15, 16, 17, 24, 25, 26}, {2, 5, 8, 11, 14, 17, 20, 23, 26}, {0, 3, 6, 9, 12, 15, 18, 21, 24}};
    return f[face];
  }
  inline void nurbs27_face_directions(int face, int* dir)
  {
    dir[0] = face < 4 ? 0 : 1;
    dir[1] = face < 2 ? 1 : 2;
  }
}  // namespace fc_oracle
