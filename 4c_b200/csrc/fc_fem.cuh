// Shape-function derivatives, Gauss rules and small dense helpers shared by host setup code and
// device kernels (all __host__ __device__).  Definitions follow the reference (paths relative
// to the reference tree):
//   shape functions  src/core/fem/src/general/utils/4C_fem_general_utils_fem_shapefunctions.hpp
//                    hex8 :53-72/:386-427, hex27 :190-230/:683-792, tet10 :244-259/:808-850,
//                    hex20 :138-189/:584-682, tet4 :231-243/:790-807, wedge6 :260-271/:854-882, pyramid5 :312-345/:953-996
//   Gauss rules      src/core/fem/src/general/utils/4C_fem_general_utils_integration.cpp
//                    hex_8point :74-105, hex_27point :130-243, tet_1point :322-330, tet_4point :331-350,
//                    wedge_6point :5992-6030, pyramid_8point :6092-6133
//   rule selection   src/solid_3D_ele/4C_solid_3D_ele_calc_lib_integration.hpp:39-62 and
//                    src/core/fem/src/general/element/4C_fem_general_element_integration_select.hpp:49-82
//   node parameter coordinates (the points of the node-Jacobian check)
//                    src/core/fem/src/general/utils/4C_fem_general_utils_local_connectivity_matrices.hpp:291-327, :564-575
//                    (wedge6 takes the first six points of the wedge18 table: they are not where its own shape functions are 1)
#pragma once
#include <cmath>

#if defined(__CUDACC__)
#define FC_HD __host__ __device__ __forceinline__
#else
#define FC_HD inline
#endif

namespace fc
{
  enum
  {
    HEX8 = 0,
    HEX27 = 1,
    TET10 = 2,
    HEX20 = 3,
    TET4 = 4,
    WEDGE6 = 5,
    PYRAMID5 = 6,
    HEX18 = 7,  // quad9 in (r, s) x line2 in t: nodes 0-8 at t = -1 (corners, edge mids, centre), 9-17 at t = +1
    NURBS27 = 8  // 3 x 3 x 3 control points, u fastest; shape functions per element from its knot spans and weights (fc_nurbs.cuh)
  };

  template <int CT>
  struct Cell;
  template <>
  struct Cell<HEX8>
  {
    static constexpr int nen = 8, ngp = 8;
  };
  template <>
  struct Cell<HEX27>
  {
    static constexpr int nen = 27, ngp = 27;
  };
  template <>
  struct Cell<TET10>
  {
    static constexpr int nen = 10, ngp = 4;
  };

  template <>
  struct Cell<HEX20>
  {
    static constexpr int nen = 20, ngp = 27;
  };
  template <>
  struct Cell<TET4>
  {
    static constexpr int nen = 4, ngp = 1;
  };
  template <>
  struct Cell<WEDGE6>
  {
    static constexpr int nen = 6, ngp = 6;
  };
  template <>
  struct Cell<PYRAMID5>
  {
    static constexpr int nen = 5, ngp = 8;
  };

  template <>
  struct Cell<HEX18>
  {
    static constexpr int nen = 18, ngp = 18;  // hex_18point = quad_9point x line_2point
  };

  template <>
  struct Cell<NURBS27>
  {
    static constexpr int nen = 27, ngp = 27;  // hex_27point (4C_fem_general_element_integration_select.hpp:89-92)
  };

  FC_HD int cell_nen(int ct)
  {
    switch (ct)
    {
      case HEX8: return Cell<HEX8>::nen;
      case HEX27: return Cell<HEX27>::nen;
      case TET10: return Cell<TET10>::nen;
      case HEX20: return Cell<HEX20>::nen;
      case TET4: return Cell<TET4>::nen;
      case WEDGE6: return Cell<WEDGE6>::nen;
      case HEX18: return Cell<HEX18>::nen;
      case NURBS27: return Cell<NURBS27>::nen;
      default: return Cell<PYRAMID5>::nen;
    }
  }
  FC_HD int cell_ngp(int ct)
  {
    switch (ct)
    {
      case HEX8: return Cell<HEX8>::ngp;
      case HEX27: return Cell<HEX27>::ngp;
      case TET10: return Cell<TET10>::ngp;
      case HEX20: return Cell<HEX20>::ngp;
      case TET4: return Cell<TET4>::ngp;
      case WEDGE6: return Cell<WEDGE6>::ngp;
      case HEX18: return Cell<HEX18>::ngp;
      case NURBS27: return Cell<NURBS27>::ngp;
      default: return Cell<PYRAMID5>::ngp;
    }
  }

  // Parameter-space sign/position of hex node a in direction d as an integer in {-1,0,1}.
  // Reference numbering: corners 0-7, edge mids 8-19, face centres 20-25, centre 26.
  FC_HD int hex_node_pos(int a, int d)
  {
    // packed as base-3 digits (value+1) for (r,s,t)
    const unsigned char tab[27] = {
        0 + 0 * 3 + 0 * 9, 2 + 0 * 3 + 0 * 9, 2 + 2 * 3 + 0 * 9, 0 + 2 * 3 + 0 * 9,  // 0-3
        0 + 0 * 3 + 2 * 9, 2 + 0 * 3 + 2 * 9, 2 + 2 * 3 + 2 * 9, 0 + 2 * 3 + 2 * 9,  // 4-7
        1 + 0 * 3 + 0 * 9, 2 + 1 * 3 + 0 * 9, 1 + 2 * 3 + 0 * 9, 0 + 1 * 3 + 0 * 9,  // 8-11
        0 + 0 * 3 + 1 * 9, 2 + 0 * 3 + 1 * 9, 2 + 2 * 3 + 1 * 9, 0 + 2 * 3 + 1 * 9,  // 12-15
        1 + 0 * 3 + 2 * 9, 2 + 1 * 3 + 2 * 9, 1 + 2 * 3 + 2 * 9, 0 + 1 * 3 + 2 * 9,  // 16-19
        1 + 1 * 3 + 0 * 9, 1 + 0 * 3 + 1 * 9, 2 + 1 * 3 + 1 * 9, 1 + 2 * 3 + 1 * 9,  // 20-23
        0 + 1 * 3 + 1 * 9, 1 + 1 * 3 + 2 * 9, 1 + 1 * 3 + 1 * 9};                    // 24-26
    int v = tab[a];
    if (d == 1) v /= 3;
    if (d == 2) v /= 9;
    return v % 3 - 1;
  }

  // position of node / Gauss point k of the quad9 pattern (corners 0-3, edge mids 4-7, centre 8) in direction d, in {-1,0,1}
  FC_HD int quad9_pos(int k, int d) { return k < 4 ? hex_node_pos(k, d) : (k < 8 ? hex_node_pos(k + 4, d) : 0); }

  FC_HD void node_param_coords(int ct, int a, double* xi)
  {
    if (ct == NURBS27)
    {
      // eleNodeNumbering_nurbs27_nodes_reference, 4C_fem_general_utils_local_connectivity_matrices.hpp:305-314
      xi[0] = (double)(a % 3 - 1);
      xi[1] = (double)((a / 3) % 3 - 1);
      xi[2] = (double)(a / 9 - 1);
      return;
    }
    if (ct == HEX18)
    {
      // eleNodeNumbering_hex18_nodes_reference, 4C_fem_general_utils_local_connectivity_matrices.hpp:175-180
      xi[0] = (double)quad9_pos(a % 9, 0);
      xi[1] = (double)quad9_pos(a % 9, 1);
      xi[2] = a < 9 ? -1.0 : 1.0;
      return;
    }
    if (ct == WEDGE6)
    {
      const double t[6][3] = {{0, 0, -1}, {1, 0, -1}, {0, 1, -1}, {0, 0, 1}, {1, 0, 1}, {0, 1, 1}};
      for (int d = 0; d < 3; ++d) xi[d] = t[a][d];
    }
    else if (ct == PYRAMID5)
    {
      const double t[5][3] = {{-1, -1, 0}, {1, -1, 0}, {1, 1, 0}, {-1, 1, 0}, {0, 0, 1}};
      for (int d = 0; d < 3; ++d) xi[d] = t[a][d];
    }
    else if (ct == TET10 || ct == TET4)
    {
      // corners (0,0,0),(1,0,0),(0,1,0),(0,0,1); mids 4:(01) 5:(12) 6:(02) 7:(03) 8:(13) 9:(23)
      const double t[10][3] = {{0, 0, 0}, {1, 0, 0}, {0, 1, 0}, {0, 0, 1}, {.5, 0, 0}, {.5, .5, 0},
          {0, .5, 0}, {0, 0, .5}, {.5, 0, .5}, {0, .5, .5}};
      xi[0] = t[a][0];
      xi[1] = t[a][1];
      xi[2] = t[a][2];
    }
    else
    {
      for (int d = 0; d < 3; ++d) xi[d] = (double)hex_node_pos(a, d);
    }
  }

  // dN[3*a+d] = dN_a/dxi_d
  template <int CT>
  FC_HD void shape_deriv1(const double* xi, double* dN)
  {
    static_assert(CT != NURBS27, "nurbs27 has no parameter-space shape functions of its own: element tables, fc_nurbs.cuh");
    const double r = xi[0], s = xi[1], t = xi[2];
    if constexpr (CT == HEX8)
    {
      const double q = 0.125;
      const double rf[2] = {1.0 - r, 1.0 + r}, sf[2] = {1.0 - s, 1.0 + s}, tf[2] = {1.0 - t, 1.0 + t};
#pragma unroll
      for (int a = 0; a < 8; ++a)
      {
        const int ir = hex_node_pos(a, 0) > 0, is = hex_node_pos(a, 1) > 0, it = hex_node_pos(a, 2) > 0;
        dN[3 * a + 0] = (ir ? q : -q) * sf[is] * tf[it];
        dN[3 * a + 1] = (is ? q : -q) * tf[it] * rf[ir];
        dN[3 * a + 2] = (it ? q : -q) * rf[ir] * sf[is];
      }
    }
    else if constexpr (CT == HEX18)
    {
      // 4C_fem_general_utils_fem_shapefunctions.hpp:508-560: quadratic Lagrange in r and s, linear in t
      const double R[3] = {0.5 * r * (r - 1.0), 1.0 - r * r, 0.5 * r * (r + 1.0)};
      const double S[3] = {0.5 * s * (s - 1.0), 1.0 - s * s, 0.5 * s * (s + 1.0)};
      const double dR[3] = {r - 0.5, -2.0 * r, r + 0.5};
      const double dS[3] = {s - 0.5, -2.0 * s, s + 0.5};
      for (int a = 0; a < 18; ++a)
      {
        const int ir = quad9_pos(a % 9, 0) + 1, is = quad9_pos(a % 9, 1) + 1;
        const double L = a < 9 ? 0.5 * (1.0 - t) : 0.5 * (1.0 + t);
        dN[3 * a + 0] = dR[ir] * S[is] * L;
        dN[3 * a + 1] = R[ir] * dS[is] * L;
        dN[3 * a + 2] = R[ir] * S[is] * (a < 9 ? -0.5 : 0.5);
      }
    }
    else if constexpr (CT == HEX27)
    {
      const double R[3] = {0.5 * r * (r - 1.0), 1.0 - r * r, 0.5 * r * (r + 1.0)};
      const double S[3] = {0.5 * s * (s - 1.0), 1.0 - s * s, 0.5 * s * (s + 1.0)};
      const double T[3] = {0.5 * t * (t - 1.0), 1.0 - t * t, 0.5 * t * (t + 1.0)};
      const double dR[3] = {r - 0.5, -2.0 * r, r + 0.5};
      const double dS[3] = {s - 0.5, -2.0 * s, s + 0.5};
      const double dT[3] = {t - 0.5, -2.0 * t, t + 0.5};
      for (int a = 0; a < 27; ++a)
      {
        const int ir = hex_node_pos(a, 0) + 1, is = hex_node_pos(a, 1) + 1, it = hex_node_pos(a, 2) + 1;
        dN[3 * a + 0] = S[is] * T[it] * dR[ir];
        dN[3 * a + 1] = R[ir] * T[it] * dS[is];
        dN[3 * a + 2] = R[ir] * S[is] * dT[it];
      }
    }
    else if constexpr (CT == HEX20)
    {
      // corner a: N = 1/8 R S T (R + S + T - 5) with R = 1 + r_a r, ...; mid-edge node on an edge along r: 1/4 (1 - r^2) S T
      const double P[3] = {r, s, t};
      for (int a = 0; a < 20; ++a)
      {
        int p[3];
        double L[3];
        for (int d = 0; d < 3; ++d)
        {
          p[d] = hex_node_pos(a, d);
          L[d] = p[d] == 0 ? 1.0 - P[d] * P[d] : 1.0 + p[d] * P[d];
        }
        if (a < 8)
        {
          const double sum = L[0] + L[1] + L[2] - 5.0;
          dN[3 * a + 0] = 0.125 * p[0] * L[1] * L[2] * (sum + L[0]);
          dN[3 * a + 1] = 0.125 * p[1] * L[2] * L[0] * (sum + L[1]);
          dN[3 * a + 2] = 0.125 * p[2] * L[0] * L[1] * (sum + L[2]);
        }
        else
          for (int d = 0; d < 3; ++d)
          {
            const int d1 = (d + 1) % 3, d2 = (d + 2) % 3;
            dN[3 * a + d] = p[d] == 0 ? -0.5 * P[d] * L[d1] * L[d2] : 0.25 * p[d] * L[d1] * L[d2];
          }
      }
    }
    else if constexpr (CT == TET4)
    {
      const double v[12] = {-1, -1, -1, 1, 0, 0, 0, 1, 0, 0, 0, 1};
      for (int i = 0; i < 12; ++i) dN[i] = v[i];
    }
    else if constexpr (CT == WEDGE6)
    {
      const double p1 = 0.5 * (1.0 - t), p2 = 0.5 * (1.0 + t), t3 = 1.0 - r - s;
      const double v[18] = {p1, 0, -0.5 * r, 0, p1, -0.5 * s, -p1, -p1, -0.5 * t3, p2, 0, 0.5 * r, 0, p2, 0.5 * s, -p2, -p2, 0.5 * t3};
      for (int i = 0; i < 18; ++i) dN[i] = v[i];
    }
    else if constexpr (CT == PYRAMID5)
    {
      // the reference's rational pyramid functions; at the apex (|t - 1| <= 1e-14) it sets the three ratios to 1
      const bool apex = !(fabs(t - 1.0) > 1.0e-14);
      const double dr = apex ? 1.0 : s * t / (1.0 - t), ds = apex ? 1.0 : r * t / (1.0 - t), dt = apex ? 1.0 : r * s / ((1.0 - t) * (1.0 - t));
      const double v[15] = {0.25 * (-(1.0 - s) + dr), 0.25 * (-(1.0 - r) + ds), 0.25 * (dt - 1.0),  //
          0.25 * ((1.0 - s) - dr), 0.25 * (-(1.0 + r) - ds), 0.25 * (-dt - 1.0),                     //
          0.25 * ((1.0 + s) + dr), 0.25 * ((1.0 + r) + ds), 0.25 * (dt - 1.0),                       //
          0.25 * (-(1.0 + s) - dr), 0.25 * ((1.0 - r) - ds), 0.25 * (-dt - 1.0),                     //
          0.0, 0.0, 1.0};
      for (int i = 0; i < 15; ++i) dN[i] = v[i];
    }
    else
    {
      const double u = 1.0 - r - s - t;
      const double d0 = -4.0 * u + 1.0;
      const double v[30] = {d0, d0, d0, 4.0 * r - 1.0, 0, 0, 0, 4.0 * s - 1.0, 0, 0, 0, 4.0 * t - 1.0,
          4.0 * (u - r), -4.0 * r, -4.0 * r, 4.0 * s, 4.0 * r, 0, -4.0 * s, 4.0 * (u - s), -4.0 * s,
          -4.0 * t, -4.0 * t, 4.0 * (u - t), 4.0 * t, 0, 4.0 * r, 0, 4.0 * t, 4.0 * s};
      for (int i = 0; i < 30; ++i) dN[i] = v[i];
    }
  }

  // N[a] (4C_fem_general_utils_fem_shapefunctions.hpp:53-72 hex8, :190-230 hex27, :244-259 tet10)
  template <int CT>
  FC_HD void shape_function(const double* xi, double* N)
  {
    static_assert(CT != NURBS27, "nurbs27 has no parameter-space shape functions of its own: element tables, fc_nurbs.cuh");
    const double r = xi[0], s = xi[1], t = xi[2];
    if constexpr (CT == HEX8)
    {
      const double rf[2] = {1.0 - r, 1.0 + r}, sf[2] = {1.0 - s, 1.0 + s}, tf[2] = {1.0 - t, 1.0 + t};
      for (int a = 0; a < 8; ++a)
        N[a] = 0.125 * rf[hex_node_pos(a, 0) > 0] * sf[hex_node_pos(a, 1) > 0] * tf[hex_node_pos(a, 2) > 0];
    }
    else if constexpr (CT == HEX18)
    {
      // 4C_fem_general_utils_fem_shapefunctions.hpp:104-137
      const double R[3] = {0.5 * r * (r - 1.0), 1.0 - r * r, 0.5 * r * (r + 1.0)};
      const double S[3] = {0.5 * s * (s - 1.0), 1.0 - s * s, 0.5 * s * (s + 1.0)};
      for (int a = 0; a < 18; ++a) N[a] = R[quad9_pos(a % 9, 0) + 1] * S[quad9_pos(a % 9, 1) + 1] * (a < 9 ? 0.5 * (1.0 - t) : 0.5 * (1.0 + t));
    }
    else if constexpr (CT == HEX27)
    {
      const double R[3] = {0.5 * r * (r - 1.0), 1.0 - r * r, 0.5 * r * (r + 1.0)};
      const double S[3] = {0.5 * s * (s - 1.0), 1.0 - s * s, 0.5 * s * (s + 1.0)};
      const double T[3] = {0.5 * t * (t - 1.0), 1.0 - t * t, 0.5 * t * (t + 1.0)};
      for (int a = 0; a < 27; ++a) N[a] = R[hex_node_pos(a, 0) + 1] * S[hex_node_pos(a, 1) + 1] * T[hex_node_pos(a, 2) + 1];
    }
    else if constexpr (CT == HEX20)
    {
      const double P[3] = {r, s, t};
      for (int a = 0; a < 20; ++a)
      {
        double L[3];
        for (int d = 0; d < 3; ++d)
        {
          const int p = hex_node_pos(a, d);
          L[d] = p == 0 ? 1.0 - P[d] * P[d] : 1.0 + p * P[d];
        }
        N[a] = a < 8 ? 0.125 * L[0] * L[1] * L[2] * (L[0] + L[1] + L[2] - 5.0) : 0.25 * L[0] * L[1] * L[2];
      }
    }
    else if constexpr (CT == TET4)
    {
      N[0] = 1.0 - r - s - t;
      N[1] = r;
      N[2] = s;
      N[3] = t;
    }
    else if constexpr (CT == WEDGE6)
    {
      const double t3 = 1.0 - r - s, lo = 0.5 * (1.0 - t), hi = 0.5 * (1.0 + t);
      const double v[6] = {r * lo, s * lo, t3 * lo, r * hi, s * hi, t3 * hi};
      for (int a = 0; a < 6; ++a) N[a] = v[a];
    }
    else if constexpr (CT == PYRAMID5)
    {
      const double ration = fabs(t - 1.0) > 1.0e-14 ? (r * s * t) / (1.0 - t) : 0.0;
      N[0] = 0.25 * ((1.0 - r) * (1.0 - s) - t + ration);
      N[1] = 0.25 * ((1.0 + r) * (1.0 - s) - t - ration);
      N[2] = 0.25 * ((1.0 + r) * (1.0 + s) - t + ration);
      N[3] = 0.25 * ((1.0 - r) * (1.0 + s) - t - ration);
      N[4] = t;
    }
    else
    {
      const double u = 1.0 - r - s - t;
      const double v[10] = {u * (2.0 * u - 1.0), r * (2.0 * r - 1.0), s * (2.0 * s - 1.0), t * (2.0 * t - 1.0), 4.0 * r * u,
          4.0 * r * s, 4.0 * s * u, 4.0 * t * u, 4.0 * r * t, 4.0 * s * t};
      for (int a = 0; a < 10; ++a) N[a] = v[a];
    }
  }

  // Gauss rule of the mass integral (get_gauss_rule_mass_matrix, 4C_solid_3D_ele_calc_lib_integration.hpp:40-44):
  // the stiffness rule except for the tets: tet_11point for tet10 (4C_fem_general_utils_integration.cpp:447-495),
  // tet_4point for tet4
  template <int CT>
  struct MassRule
  {
    static constexpr int ngp = CT == TET10 ? 11 : (CT == TET4 ? 4 : Cell<CT>::ngp);
  };

  // Gauss rule of the stiffness/force integral: point g of cell type CT
  template <int CT>
  FC_HD void gauss_point(int g, double* xi, double* w)
  {
    if constexpr (CT == HEX8)
    {
      const double p = 0.57735026918962576451;  // 1/sqrt(3)
      for (int d = 0; d < 3; ++d) xi[d] = hex_node_pos(g, d) > 0 ? p : -p;
      *w = 1.0;
    }
    else if constexpr (CT == TET4)
    {
      xi[0] = xi[1] = xi[2] = 0.25;
      *w = 1.0 / 6.0;
    }
    else if constexpr (CT == WEDGE6)
    {
      const double q23 = 2.0 / 3.0, q16 = 1.0 / 6.0, p = 0.57735026918962576451;
      const int k = g % 3;
      xi[0] = k == 0 ? q23 : q16;
      xi[1] = k == 1 ? q23 : q16;
      xi[2] = g < 3 ? p : -p;
      *w = q16;
    }
    else if constexpr (CT == PYRAMID5)
    {
      // the reference's 14-digit table
      const double a = 0.26318405556971, b = 0.50661630334979;
      const double m = (g & 1) ? b : a;
      xi[0] = g < 4 ? -m : m;
      xi[1] = (g & 2) ? m : -m;
      xi[2] = (g & 1) ? 0.12251482265544 : 0.54415184401122;
      *w = (g & 1) ? 0.23254745125351 : 0.10078588207983;
    }
    else if constexpr (CT == HEX18)
    {
      // hex_18point (4C_fem_general_utils_integration.cpp:112-130): quad_9point in the plane (:6185-6233, the order of the quad9
      // nodes) times line_2point through the thickness (:6946-6953), with the reference's truncated constants
      const double p = 0.7745966692415, w1 = 0.5555555555556, w2 = 0.8888888888889;
      const int k = g % 9, sr = quad9_pos(k, 0), ss = quad9_pos(k, 1);
      xi[0] = sr == 0 ? 0.0 : (sr > 0 ? p : -p);
      xi[1] = ss == 0 ? 0.0 : (ss > 0 ? p : -p);
      xi[2] = g < 9 ? -0.5773502691896 : 0.5773502691896;
      *w = (sr == 0 ? w2 : w1) * (ss == 0 ? w2 : w1);
    }
    else if constexpr (CT == HEX27 || CT == HEX20 || CT == NURBS27)
    {
      // the reference's truncated 13-digit constants (integration.cpp:133,214-215)
      const double p = 0.7745966692415, w1 = 0.5555555555556, w2 = 0.8888888888889;
      double ww = 1.0;
      for (int d = 0; d < 3; ++d)
      {
        const int s = hex_node_pos(g, d);
        xi[d] = s == 0 ? 0.0 : (s > 0 ? p : -p);
        ww *= (s == 0 ? w2 : w1);
      }
      // reference multiplies the three 1D weights left to right in (r,s,t) order
      *w = ww;
    }
    else
    {
      const double pa = (5.0 + 3.0 * 2.2360679774997896964) / 20.0;
      const double pb = (5.0 - 2.2360679774997896964) / 20.0;
      xi[0] = xi[1] = xi[2] = pb;
      if (g > 0) xi[g - 1] = pa;
      *w = 1.0 / 24.0;
    }
  }

  template <int CT>
  FC_HD void mass_gauss_point(int g, double* xi, double* w)
  {
    if constexpr (CT == TET4)
      gauss_point<TET10>(g, xi, w);  // tet_4point
    else if constexpr (CT != TET10)
      gauss_point<CT>(g, xi, w);
    else
    {
      const double a = 0.2500000000000000, b = 0.7857142857142857, c = 0.0714285714285714;
      const double d = 0.1005964238332008, e = 0.3994035761667992;
      const double p[11][3] = {{a, a, a}, {b, c, c}, {c, c, c}, {c, c, b}, {c, b, c}, {d, e, e}, {e, d, e}, {e, e, d},
          {e, d, d}, {d, e, d}, {d, d, e}};
      for (int k = 0; k < 3; ++k) xi[k] = p[g][k];
      *w = (g == 0 ? -0.0789333333333333 : (g < 5 ? 0.0457333333333333 : 0.1493333333333333)) * (1.0 / 6.0);
    }
  }

  FC_HD void centroid_param(int ct, double* xi)
  {
    // evaluate_parameter_coordinate_centroid, 4C_solid_3D_ele_calc_lib.hpp:230-294
    const double c = (ct == TET10 || ct == TET4) ? 0.25 : 0.0;
    xi[0] = xi[1] = xi[2] = c;
    if (ct == WEDGE6) xi[0] = xi[1] = 1.0 / 3.0;
    if (ct == PYRAMID5) xi[2] = 0.25;
  }

  // ---- 3x3 helpers, row-major a[3*r+c] ------------------------------------------------------------
  FC_HD double det3(const double* a)
  {
    return a[0] * (a[4] * a[8] - a[5] * a[7]) - a[1] * (a[3] * a[8] - a[5] * a[6]) +
           a[2] * (a[3] * a[7] - a[4] * a[6]);
  }
  // inv = a^-1, returns det
  FC_HD double inv3(const double* a, double* inv)
  {
    const double c0 = a[4] * a[8] - a[5] * a[7];
    const double c1 = a[5] * a[6] - a[3] * a[8];
    const double c2 = a[3] * a[7] - a[4] * a[6];
    const double det = a[0] * c0 + a[1] * c1 + a[2] * c2;
    const double id = 1.0 / det;
    inv[0] = c0 * id;
    inv[1] = (a[2] * a[7] - a[1] * a[8]) * id;
    inv[2] = (a[1] * a[5] - a[2] * a[4]) * id;
    inv[3] = c1 * id;
    inv[4] = (a[0] * a[8] - a[2] * a[6]) * id;
    inv[5] = (a[2] * a[3] - a[0] * a[5]) * id;
    inv[6] = c2 * id;
    inv[7] = (a[1] * a[6] - a[0] * a[7]) * id;
    inv[8] = (a[0] * a[4] - a[1] * a[3]) * id;
    return det;
  }
}  // namespace fc
