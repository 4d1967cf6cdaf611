// nurbs27 solid elements (ImplementedSolidCellTypes, src/solid_3D_ele/4C_solid_3D_ele_factory.hpp:36-38): NURBS basis of
// degree 2 from the element's knot spans and control point weights.
//   B-spline value + derivative   src/core/fem/src/general/nurbs/4C_fem_general_utils_bspline.hpp:539-760
//   3D functions / derivatives    src/core/fem/src/general/nurbs/4C_fem_general_utils_nurbs_shapefunctions.hpp:2586-2931
//   element knots, zero size      src/core/fem/src/nurbs_discretization/4C_fem_nurbs_discretization_knotvector.cpp:173-243
//   face knots                    the same file :251-420
//   how the element gets them     src/solid_3D_ele/4C_solid_3D_ele_calc_lib.hpp:112-150, 193-200, 397-404
// The reference evaluates the basis at every Gauss point of every element in every Newton iteration.  Knots and weights
// belong to the mesh, so here k_nurbs_tables evaluates it once per mesh into a per-element table
//   dN [ngp + nen + 1][nen][3]   Gauss points, the 27 points of the node-Jacobian check, centroid  (the layout of the other
//                                cells' shared table, so that the element kernels only add an element offset: ele_table)
//   dN at the Gauss points in the node-major and the fragment order of the tensor-core kernel (fc_kernels_nh_mma_gen.cuh)
//   N  [ngp][nen]                for the mass matrix and the body loads
// 80 KB per element; an evaluation reads the part its kernel needs (generic kernel 17.5 KB, tensor-core kernel 39 KB).
#pragma once
#include "fc_fem.cuh"

namespace fc
{
  struct NurbsTab
  {
    static constexpr int nen = 27, ngp = 27;
    // the Gauss-point derivatives again in the two orders k_nh_mma_gen reads them in (like fill_tables, fc_host_util.hpp):
    // node-major [nen][ngp][3] and fragment order [3][KS * MT][32 lanes], then the shape function values [ngp][nen]
    static constexpr int MT = 4, KS = 7;
    static constexpr int oNodeMajor = (ngp + nen + 1) * nen * 3;
    static constexpr int oFragment = oNodeMajor + nen * ngp * 3;
    static constexpr int oN = oFragment + 3 * KS * MT * 32;
    static constexpr int stride = oN + ngp * nen;
  };

  // shape-derivative table of element e: the mesh's table for the Lagrange cells, the element's own for nurbs27
  template <int CT>
  FC_HD const double* ele_table(const double* tab, int e)
  {
    return CT == NURBS27 ? tab + (size_t)e * NurbsTab::stride : tab;
  }

  FC_HD double nurbs_ratio(double num, double den) { return fabs(den) < 10e-9 ? 0.0 : num / den; }

  // local basis function ldofid (0..2) of degree 2 on the 6 knots kn at x: value and derivative
  FC_HD void bspline2(const double* kn, double x, int ldofid, double* val, double* der)
  {
    double b[3] = {0.0, 0.0, 0.0};
    b[2 - ldofid] = 1.0;
    for (int rr = 0; rr < 2; ++rr)
    {
      const int i = ldofid + rr;
      const double f1 = nurbs_ratio(x - kn[i], kn[i + 1] - kn[i]);
      const double f2 = nurbs_ratio(kn[i + 2] - x, kn[i + 2] - kn[i + 1]);
      b[rr] = f1 * b[rr] + f2 * b[rr + 1];
    }
    const double d1 = kn[ldofid + 2] - kn[ldofid], d2 = kn[ldofid + 3] - kn[ldofid + 1];
    *val = nurbs_ratio(x - kn[ldofid], d1) * b[0] + nurbs_ratio(kn[ldofid + 3] - x, d2) * b[1];
    *der = nurbs_ratio(2.0, d1) * b[0] - nurbs_ratio(2.0, d2) * b[1];
  }

  // N [3^DIM], dN [3^DIM][DIM] at uv in [-1, 1]^DIM; knots [DIM][6]; w = weights of the 3^DIM control points (u fastest).
  // Either output may be null.
  template <int DIM>
  FC_HD void nurbs_funct_deriv(const double* uv, const double* knots, const double* w, double* N, double* dN)
  {
    constexpr int NB = DIM == 3 ? 27 : 9;
    double Jd[3] = {1.0, 1.0, 1.0}, val[3][3], der[3][3];
    for (int d = 0; d < DIM; ++d)
    {
      const double* kn = knots + 6 * d;
      Jd[d] = (kn[3] - kn[2]) / 2.0;
      const double x = kn[2] + (uv[d] + 1.0) * Jd[d];
      for (int k = 0; k < 3; ++k) bspline2(kn, x, k, &val[d][k], &der[d][k]);
    }
    double sw = 0.0, sdw[3] = {0.0, 0.0, 0.0};
    for (int id = 0; id < NB; ++id)
    {
      const int rr = id % 3, mm = (id / 3) % 3, nn = id / 9;
      const double vn = DIM == 3 ? val[2][nn] : 1.0, dn = DIM == 3 ? der[2][nn] : 0.0;
      sw += w[id] * (val[0][rr] * val[1][mm] * vn);
      sdw[0] += w[id] * (der[0][rr] * val[1][mm] * vn);
      sdw[1] += w[id] * (val[0][rr] * der[1][mm] * vn);
      if (DIM == 3) sdw[2] += w[id] * (val[0][rr] * val[1][mm] * dn);
    }
    for (int id = 0; id < NB; ++id)
    {
      const int rr = id % 3, mm = (id / 3) % 3, nn = id / 9;
      const double vn = DIM == 3 ? val[2][nn] : 1.0, dn = DIM == 3 ? der[2][nn] : 0.0;
      const double B = val[0][rr] * val[1][mm] * vn;
      const double dB[3] = {der[0][rr] * val[1][mm] * vn, val[0][rr] * der[1][mm] * vn, val[0][rr] * val[1][mm] * dn};
      if (N) N[id] = w[id] * B / sw;
      if (dN)
        for (int d = 0; d < DIM; ++d)
        {
          double v = dB[d] * sw - B * sdw[d];
          v *= w[id] / (sw * sw);
          dN[DIM * id + d] = v * Jd[d];
        }
    }
  }

#if defined(__CUDACC__)
  // one thread per (element, point): points 0..26 Gauss points, 27..53 node-check points, 54 centroid.
  // bad[e] is set for an element with a zero-sized knot span (the reference throws, calc_lib.hpp:196-199).
  static __global__ void k_nurbs_tables(const double* __restrict__ knots, const double* __restrict__ weights, const int* __restrict__ conn,
      int n_ele, double* __restrict__ tab, unsigned char* __restrict__ zero_size)
  {
    using T = NurbsTab;
    constexpr int NP = T::ngp + T::nen + 1;
    const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (tid >= (long long)n_ele * NP) return;
    const int e = (int)(tid / NP), p = (int)(tid % NP);
    const double* kn = knots + (size_t)e * 18;
    double w[27], xi[3] = {0.0, 0.0, 0.0}, gw;
    for (int a = 0; a < 27; ++a) w[a] = weights[conn[(size_t)e * 27 + a]];
    if (p < T::ngp)
      gauss_point<NURBS27>(p, xi, &gw);
    else if (p < T::ngp + T::nen)
      node_param_coords(NURBS27, p - T::ngp, xi);
    double* out = tab + (size_t)e * T::stride;
    double* dN = out + (size_t)p * T::nen * 3;
    nurbs_funct_deriv<3>(xi, kn, w, p < T::ngp ? out + T::oN + p * T::nen : nullptr, dN);
    if (p < T::ngp)  // the padding entries of the fragment order stay zero (the table is cleared before this kernel)
      for (int a = 0; a < T::nen; ++a)
        for (int d = 0; d < 3; ++d)
        {
          out[T::oNodeMajor + (a * T::ngp + p) * 3 + d] = dN[3 * a + d];
          out[T::oFragment + (d * T::KS * T::MT + (p >> 2) * T::MT + (a >> 3)) * 32 + ((a & 7) << 2 | (p & 3))] = dN[3 * a + d];
        }
    if (p == 0)
    {
      bool zs = false;
      for (int d = 0; d < 3; ++d)
        if (fabs(kn[6 * d + 2] - kn[6 * d + 3]) < 1e-12) zs = true;
      zero_size[e] = zs ? 1 : 0;
    }
  }
#endif
}  // namespace fc
