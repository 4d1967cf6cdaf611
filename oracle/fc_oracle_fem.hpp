// ORACLE — TEST INFRASTRUCTURE ONLY. Never linked into or called by the product library.
//
// Shape functions, their first derivatives, node parameter coordinates and Gauss rules for
// hex8 / hex27 / tet10 as the reference defines them (paths relative to /root/reference):
//   shape functions   src/core/fem/src/general/utils/4C_fem_general_utils_fem_shapefunctions.hpp
//                     hex8 :53-72 / :386-427, hex27 :190-230 / :683-792, tet10 :244-259 / :808-850
//   Gauss rules       src/core/fem/src/general/utils/4C_fem_general_utils_integration.cpp
//                     hex_8point :74-105, hex_27point :130-243 (13-digit constants!), tet_4point :331-350
//   rule selection    src/solid_3D_ele/4C_solid_3D_ele_calc_lib_integration.hpp:39-62
//                     (tet10 stiffness uses the 4-point rule)
#pragma once
#include <cmath>

namespace fc_oracle
{
  enum CellType
  {
    HEX8 = 0,
    HEX27 = 1,
    TET10 = 2
  };

  inline int num_nodes(int ct) { return ct == HEX8 ? 8 : (ct == HEX27 ? 27 : 10); }
  inline int num_gp(int ct) { return ct == HEX8 ? 8 : (ct == HEX27 ? 27 : 4); }

  // parameter coordinates (r,s,t) of the element nodes, reference node numbering.
  // hex8/hex27: corner nodes, then edge mid nodes 8..19, face centres 20..25, centre 26.
  inline void node_param_coords(int ct, int a, double* xi)
  {
    static const signed char hex[27][3] = {{-1, -1, -1}, {1, -1, -1}, {1, 1, -1}, {-1, 1, -1},
        {-1, -1, 1}, {1, -1, 1}, {1, 1, 1}, {-1, 1, 1}, {0, -1, -1}, {1, 0, -1}, {0, 1, -1},
        {-1, 0, -1}, {-1, -1, 0}, {1, -1, 0}, {1, 1, 0}, {-1, 1, 0}, {0, -1, 1}, {1, 0, 1},
        {0, 1, 1}, {-1, 0, 1}, {0, 0, -1}, {0, -1, 0}, {1, 0, 0}, {0, 1, 0}, {-1, 0, 0},
        {0, 0, 1}, {0, 0, 0}};
    static const double tet[10][3] = {{0, 0, 0}, {1, 0, 0}, {0, 1, 0}, {0, 0, 1}, {.5, 0, 0},
        {.5, .5, 0}, {0, .5, 0}, {0, 0, .5}, {.5, 0, .5}, {0, .5, .5}};
    if (ct == TET10)
      for (int d = 0; d < 3; ++d) xi[d] = tet[a][d];
    else
      for (int d = 0; d < 3; ++d) xi[d] = hex[a][d];
  }

  // N_a(xi)
  inline void shape_function(int ct, const double* xi, double* N)
  {
    const double r = xi[0], s = xi[1], t = xi[2];
    if (ct == HEX8)
    {
      const double Q18 = 1.0 / 8.0;
      const double rp = 1.0 + r, rm = 1.0 - r, sp = 1.0 + s, sm = 1.0 - s, tp = 1.0 + t,
                   tm = 1.0 - t;
      N[0] = Q18 * rm * sm * tm;
      N[1] = Q18 * rp * sm * tm;
      N[2] = Q18 * rp * sp * tm;
      N[3] = Q18 * rm * sp * tm;
      N[4] = Q18 * rm * sm * tp;
      N[5] = Q18 * rp * sm * tp;
      N[6] = Q18 * rp * sp * tp;
      N[7] = Q18 * rm * sp * tp;
    }
    else if (ct == HEX27)
    {
      // tensor-product quadratic Lagrange: index 0 -> (-1), 1 -> (0), 2 -> (+1)
      const double R[3] = {0.5 * r * (r - 1.0), (1.0 - r * r), 0.5 * r * (r + 1.0)};
      const double S[3] = {0.5 * s * (s - 1.0), (1.0 - s * s), 0.5 * s * (s + 1.0)};
      const double T[3] = {0.5 * t * (t - 1.0), (1.0 - t * t), 0.5 * t * (t + 1.0)};
      for (int a = 0; a < 27; ++a)
      {
        double x[3];
        node_param_coords(HEX27, a, x);
        N[a] = R[(int)x[0] + 1] * S[(int)x[1] + 1] * T[(int)x[2] + 1];
      }
    }
    else
    {
      const double u = 1.0 - r - s - t;
      N[0] = u * (2.0 * u - 1.0);
      N[1] = r * (2.0 * r - 1.0);
      N[2] = s * (2.0 * s - 1.0);
      N[3] = t * (2.0 * t - 1.0);
      N[4] = 4.0 * r * u;
      N[5] = 4.0 * r * s;
      N[6] = 4.0 * s * u;
      N[7] = 4.0 * t * u;
      N[8] = 4.0 * r * t;
      N[9] = 4.0 * s * t;
    }
  }

  // dN(d, a) = dN_a/dxi_d, stored column-major 3 x nen  (dN[d + 3*a])
  inline void shape_function_deriv1(int ct, const double* xi, double* dN)
  {
    const double r = xi[0], s = xi[1], t = xi[2];
    if (ct == HEX8)
    {
      const double Q18 = 1.0 / 8.0;
      const double rp = 1.0 + r, rm = 1.0 - r, sp = 1.0 + s, sm = 1.0 - s, tp = 1.0 + t,
                   tm = 1.0 - t;
      // sign pattern of node a in direction d equals its parameter coordinate
      const double Rf[2] = {rm, rp}, Sf[2] = {sm, sp}, Tf[2] = {tm, tp};
      for (int a = 0; a < 8; ++a)
      {
        double x[3];
        node_param_coords(HEX8, a, x);
        const int ir = x[0] > 0, is = x[1] > 0, it = x[2] > 0;
        // reference factor order: d/dr: Q18*s*t ; d/ds: Q18*t*r ; d/dt: Q18*r*s
        dN[0 + 3 * a] = (x[0] > 0 ? Q18 : -Q18) * Sf[is] * Tf[it];
        dN[1 + 3 * a] = (x[1] > 0 ? Q18 : -Q18) * Tf[it] * Rf[ir];
        dN[2 + 3 * a] = (x[2] > 0 ? Q18 : -Q18) * Rf[ir] * Sf[is];
      }
    }
    else if (ct == HEX27)
    {
      const double R[3] = {0.5 * r * (r - 1.0), (1.0 - r * r), 0.5 * r * (r + 1.0)};
      const double S[3] = {0.5 * s * (s - 1.0), (1.0 - s * s), 0.5 * s * (s + 1.0)};
      const double T[3] = {0.5 * t * (t - 1.0), (1.0 - t * t), 0.5 * t * (t + 1.0)};
      const double dR[3] = {r - 0.5, -2.0 * r, r + 0.5};
      const double dS[3] = {s - 0.5, -2.0 * s, s + 0.5};
      const double dT[3] = {t - 0.5, -2.0 * t, t + 0.5};
      for (int a = 0; a < 27; ++a)
      {
        double x[3];
        node_param_coords(HEX27, a, x);
        const int ir = (int)x[0] + 1, is = (int)x[1] + 1, it = (int)x[2] + 1;
        dN[0 + 3 * a] = S[is] * T[it] * dR[ir];
        dN[1 + 3 * a] = R[ir] * T[it] * dS[is];
        dN[2 + 3 * a] = R[ir] * S[is] * dT[it];
      }
    }
    else
    {
      const double u = 1.0 - r - s - t;
      const double d0 = -4.0 * u + 1.;
      const double v[10][3] = {{d0, d0, d0}, {4.0 * r - 1., 0, 0}, {0, 4.0 * s - 1., 0},
          {0, 0, 4.0 * t - 1.}, {4.0 * (u - r), -4.0 * r, -4.0 * r}, {4.0 * s, 4.0 * r, 0},
          {-4.0 * s, 4.0 * (u - s), -4.0 * s}, {-4.0 * t, -4.0 * t, 4.0 * (u - t)},
          {4.0 * t, 0, 4.0 * r}, {0, 4.0 * t, 4.0 * s}};
      for (int a = 0; a < 10; ++a)
        for (int d = 0; d < 3; ++d) dN[d + 3 * a] = v[a][d];
    }
  }

  // Gauss rule used for the stiffness/force integral of a cell type.
  inline void gauss_rule(int ct, double (*xi)[3], double* w)
  {
    if (ct == HEX8)
    {
      const double xi2 = 1.0 / std::sqrt(3.0);
      for (int g = 0; g < 8; ++g)
      {
        double x[3];
        node_param_coords(HEX8, g, x);  // points ordered like the corner nodes
        for (int d = 0; d < 3; ++d) xi[g][d] = x[d] > 0 ? xi2 : -xi2;
        w[g] = 1.0;
      }
    }
    else if (ct == HEX27)
    {
      // the reference hard-codes truncated constants (integration.cpp:133,214-215)
      const double xi3 = 0.7745966692415;
      const double w1 = 0.5555555555556;
      const double w2 = 0.8888888888889;
      for (int g = 0; g < 27; ++g)
      {
        double x[3];
        node_param_coords(HEX27, g, x);  // points ordered like the 27 nodes
        double wd[3];
        for (int d = 0; d < 3; ++d)
        {
          xi[g][d] = x[d] == 0 ? 0.0 : (x[d] > 0 ? xi3 : -xi3);
          wd[d] = x[d] == 0 ? w2 : w1;
        }
        w[g] = wd[0] * wd[1] * wd[2];
      }
    }
    else
    {
      const double palpha = (5.0 + 3.0 * std::sqrt(5.0)) / 20.0;
      const double pbeta = (5.0 - std::sqrt(5.0)) / 20.0;
      for (int g = 0; g < 4; ++g)
      {
        for (int d = 0; d < 3; ++d) xi[g][d] = pbeta;
        if (g > 0) xi[g][g - 1] = palpha;
        w[g] = 1.0 / 24.0;
      }
    }
  }

  // Gauss rule of the mass matrix where it differs from the stiffness rule: tet10 -> tet_11point
  // (4C_solid_3D_ele_calc_lib_integration.hpp:40-44 + 4C_fem_general_element_integration_select.hpp:64-67;
  //  table 4C_fem_general_utils_integration.cpp:447-495, Q16 = 1/6).  Returns the number of points.
  inline int gauss_rule_mass_tet10(double (*xi)[3], double* w)
  {
    const double a = 0.2500000000000000, b = 0.7857142857142857, c = 0.0714285714285714;
    const double d = 0.1005964238332008, e = 0.3994035761667992;
    const double p[11][3] = {{a, a, a}, {b, c, c}, {c, c, c}, {c, c, b}, {c, b, c}, {d, e, e}, {e, d, e}, {e, e, d}, {e, d, d},
        {d, e, d}, {d, d, e}};
    const double Q16 = 1.0 / 6.0;
    for (int g = 0; g < 11; ++g)
    {
      for (int k = 0; k < 3; ++k) xi[g][k] = p[g][k];
      w[g] = (g == 0 ? -0.0789333333333333 : (g < 5 ? 0.0457333333333333 : 0.1493333333333333)) * Q16;
    }
    return 11;
  }

  // centroid in parameter space (used by F-bar): hex -> origin, tet -> (1/4,1/4,1/4)
  inline void centroid_param(int ct, double* xi)
  {
    const double c = ct == TET10 ? 0.25 : 0.0;
    xi[0] = xi[1] = xi[2] = c;
  }
}  // namespace fc_oracle
