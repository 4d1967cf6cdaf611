// ORACLE — TEST INFRASTRUCTURE ONLY. Never linked into or called by the product library.
//
// Shape functions, their first derivatives, node parameter coordinates and Gauss rules for
// hex8 / hex27 / tet10 / hex20 / tet4 / wedge6 / pyramid5 as the reference defines them (paths relative to /root/reference):
//   shape functions   src/core/fem/src/general/utils/4C_fem_general_utils_fem_shapefunctions.hpp
//                     hex8 :53-72 / :386-427, hex27 :190-230 / :683-792, tet10 :244-259 / :808-850,
//                     hex20 :138-189 / :584-682, tet4 :231-243 / :790-807, wedge6 :260-271 / :854-882,
//                     pyramid5 :312-345 / :953-996 (PY5TOL = 1e-14, :18)
//   Gauss rules       src/core/fem/src/general/utils/4C_fem_general_utils_integration.cpp
//                     hex_8point :74-105, hex_27point :130-243 (13-digit constants!), tet_1point :322-330,
//                     tet_4point :331-350, wedge_6point :5992-6030, pyramid_8point :6092-6133 (14-digit constants)
//   rule selection    src/solid_3D_ele/4C_solid_3D_ele_calc_lib_integration.hpp:39-62
//                     (tet10 stiffness uses the 4-point rule, tet4 stiffness the 1-point rule) over
//                     src/core/fem/src/general/element/4C_fem_general_element_integration_select.hpp:49-82
//   node parameter coordinates  src/core/fem/src/general/utils/4C_fem_general_utils_local_connectivity_matrices.hpp
//                     :291-327 (tables), :564-575 (wedge6 reads the first six rows of the wedge18 table)
#pragma once
#include <cmath>

#include "fc_oracle_nurbs.hpp"

namespace fc_oracle
{
  enum CellType
  {
    HEX8 = 0,
    HEX27 = 1,
    TET10 = 2,
    HEX20 = 3,
    TET4 = 4,
    WEDGE6 = 5,
    PYRAMID5 = 6,
    HEX18 = 7,
    NURBS27 = 8  // needs the element's knots and weights: nurbs_context(), fc_oracle_nurbs.hpp
  };

  inline int num_nodes(int ct)
  {
    static const int n[9] = {8, 27, 10, 20, 4, 6, 5, 18, 27};
    return n[ct];
  }
  inline int num_gp(int ct)
  {
    static const int n[9] = {8, 27, 4, 27, 1, 6, 8, 18, 27};
    return n[ct];
  }

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
    static const double wedge[6][3] = {{0, 0, -1}, {1, 0, -1}, {0, 1, -1}, {0, 0, 1}, {1, 0, 1}, {0, 1, 1}};
    static const double pyramid[5][3] = {{-1, -1, 0}, {1, -1, 0}, {1, 1, 0}, {-1, 1, 0}, {0, 0, 1}};
    // hex18 (4C_fem_general_utils_local_connectivity_matrices.hpp:175-180): the quad9 pattern at t = -1, then at t = +1
    static const signed char q9[9][2] = {{-1, -1}, {1, -1}, {1, 1}, {-1, 1}, {0, -1}, {1, 0}, {0, 1}, {-1, 0}, {0, 0}};
    if (ct == NURBS27)
      nurbs27_node_param(a, xi);
    else if (ct == HEX18)
    {
      xi[0] = q9[a % 9][0];
      xi[1] = q9[a % 9][1];
      xi[2] = a < 9 ? -1.0 : 1.0;
    }
    else if (ct == TET10 || ct == TET4)
      for (int d = 0; d < 3; ++d) xi[d] = tet[a][d];
    else if (ct == WEDGE6)
      for (int d = 0; d < 3; ++d) xi[d] = wedge[a][d];
    else if (ct == PYRAMID5)
      for (int d = 0; d < 3; ++d) xi[d] = pyramid[a][d];
    else
      for (int d = 0; d < 3; ++d) xi[d] = hex[a][d];
  }

  // hex18: the nine in-plane functions of (r, s) and their derivatives in the factored form of the reference
  // (4C_fem_general_utils_fem_shapefunctions.hpp:104-137, 508-560); node order: corners, edge mids, centre
  inline void hex18_inplane(double r, double s, double* q, double* qr, double* qs)
  {
    const double rp = 1.0 + r, rm = 1.0 - r, sp = 1.0 + s, sm = 1.0 - s;
    const double r2 = 1.0 - r * r, s2 = 1.0 - s * s, rh = 0.5 * r, sh = 0.5 * s, rs = rh * sh;
    const double rhp = r + 0.5, rhm = r - 0.5, shp = s + 0.5, shm = s - 0.5;
    q[0] = rs * rm * sm;   qr[0] = -rhm * sh * sm;     qs[0] = -shm * rh * rm;
    q[1] = -rs * rp * sm;  qr[1] = -rhp * sh * sm;     qs[1] = shm * rh * rp;
    q[2] = rs * rp * sp;   qr[2] = rhp * sh * sp;      qs[2] = shp * rh * rp;
    q[3] = -rs * rm * sp;  qr[3] = rhm * sh * sp;      qs[3] = -shp * rh * rm;
    q[4] = -sh * sm * r2;  qr[4] = 2.0 * r * sh * sm;  qs[4] = shm * r2;
    q[5] = rh * rp * s2;   qr[5] = rhp * s2;           qs[5] = -2.0 * s * rh * rp;
    q[6] = sh * sp * r2;   qr[6] = -2.0 * r * sh * sp; qs[6] = shp * r2;
    q[7] = -rh * rm * s2;  qr[7] = rhm * s2;           qs[7] = 2.0 * s * rh * rm;
    q[8] = r2 * s2;        qr[8] = -2.0 * r * s2;      qs[8] = -2.0 * s * r2;
  }

  // N_a(xi)
  inline void shape_function(int ct, const double* xi, double* N)
  {
    const double r = xi[0], s = xi[1], t = xi[2];
    if (ct == NURBS27)
      nurbs_funct_deriv<3>(xi, nurbs_context().knots, nurbs_context().weights, N, nullptr);
    else if (ct == HEX8)
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
    else if (ct == HEX18)
    {
      // 4C_fem_general_utils_fem_shapefunctions.hpp:104-137: nine in-plane functions times (1 -+ t) / 2
      double q[9], qr[9], qs[9];
      hex18_inplane(r, s, q, qr, qs);
      for (int k = 0; k < 9; ++k)
      {
        N[k] = q[k] * (1.0 - t) * 0.5;
        N[9 + k] = q[k] * (1.0 + t) * 0.5;
      }
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
    else if (ct == HEX20)
    {
      const double Q18 = 1.0 / 8.0;
      const double rp = 1.0 + r, rm = 1.0 - r, sp = 1.0 + s, sm = 1.0 - s, tp = 1.0 + t, tm = 1.0 - t;
      const double rrm = 1.0 - r * r, ssm = 1.0 - s * s, ttm = 1.0 - t * t;
      N[0] = Q18 * rm * sm * tm * (rm + sm + tm - 5.0);
      N[1] = Q18 * rp * sm * tm * (rp + sm + tm - 5.0);
      N[2] = Q18 * rp * sp * tm * (rp + sp + tm - 5.0);
      N[3] = Q18 * rm * sp * tm * (rm + sp + tm - 5.0);
      N[4] = Q18 * rm * sm * tp * (rm + sm + tp - 5.0);
      N[5] = Q18 * rp * sm * tp * (rp + sm + tp - 5.0);
      N[6] = Q18 * rp * sp * tp * (rp + sp + tp - 5.0);
      N[7] = Q18 * rm * sp * tp * (rm + sp + tp - 5.0);
      N[8] = 0.25 * rrm * sm * tm;
      N[9] = 0.25 * rp * ssm * tm;
      N[10] = 0.25 * rrm * sp * tm;
      N[11] = 0.25 * rm * ssm * tm;
      N[12] = 0.25 * rm * sm * ttm;
      N[13] = 0.25 * rp * sm * ttm;
      N[14] = 0.25 * rp * sp * ttm;
      N[15] = 0.25 * rm * sp * ttm;
      N[16] = 0.25 * rrm * sm * tp;
      N[17] = 0.25 * rp * ssm * tp;
      N[18] = 0.25 * rrm * sp * tp;
      N[19] = 0.25 * rm * ssm * tp;
    }
    else if (ct == TET4)
    {
      N[0] = 1.0 - r - s - t;
      N[1] = r;
      N[2] = s;
      N[3] = t;
    }
    else if (ct == WEDGE6)
    {
      const double Q12 = 1.0 / 2.0, t3 = 1.0 - r - s;
      N[0] = Q12 * r * (1.0 - t);
      N[1] = Q12 * s * (1.0 - t);
      N[2] = Q12 * t3 * (1.0 - t);
      N[3] = Q12 * r * (1.0 + t);
      N[4] = Q12 * s * (1.0 + t);
      N[5] = Q12 * t3 * (1.0 + t);
    }
    else if (ct == PYRAMID5)
    {
      const double Q14 = 1.0 / 4.0;
      double ration = 0.0;
      if (std::abs(t - 1.0) > 1.0e-14) ration = (r * s * t) / (1.0 - t);
      N[0] = Q14 * ((1.0 - r) * (1.0 - s) - t + ration);
      N[1] = Q14 * ((1.0 + r) * (1.0 - s) - t - ration);
      N[2] = Q14 * ((1.0 + r) * (1.0 + s) - t + ration);
      N[3] = Q14 * ((1.0 - r) * (1.0 + s) - t - ration);
      N[4] = t;
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
    if (ct == NURBS27)
      nurbs_funct_deriv<3>(xi, nurbs_context().knots, nurbs_context().weights, nullptr, dN);
    else if (ct == HEX8)
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
    else if (ct == HEX18)
    {
      // 4C_fem_general_utils_fem_shapefunctions.hpp:508-570
      double q[9], qr[9], qs[9];
      hex18_inplane(r, s, q, qr, qs);
      for (int k = 0; k < 9; ++k)
      {
        dN[0 + 3 * k] = qr[k] * (1.0 - t) * 0.5;
        dN[1 + 3 * k] = qs[k] * (1.0 - t) * 0.5;
        dN[2 + 3 * k] = -0.5 * q[k];
        dN[0 + 3 * (9 + k)] = qr[k] * (1.0 + t) * 0.5;
        dN[1 + 3 * (9 + k)] = qs[k] * (1.0 + t) * 0.5;
        dN[2 + 3 * (9 + k)] = 0.5 * q[k];
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
    else if (ct == HEX20)
    {
      const double Q18 = 1.0 / 8.0;
      const double rp = 1.0 + r, rm = 1.0 - r, sp = 1.0 + s, sm = 1.0 - s, tp = 1.0 + t, tm = 1.0 - t;
      const double rrm = 1.0 - r * r, ssm = 1.0 - s * s, ttm = 1.0 - t * t;
      // corner a: factors R, S, T in {m, p} by the sign of its position; the reference writes, e.g. for node 0,
      //   d/dr = -Q18 sm tm (2 rm + sm + tm - 5),  d/ds = -Q18 tm rm (2 sm + tm + rm - 5),  d/dt = -Q18 rm sm (2 tm + rm + sm - 5)
      const double Rf[2] = {rm, rp}, Sf[2] = {sm, sp}, Tf[2] = {tm, tp};
      for (int a = 0; a < 8; ++a)
      {
        double x[3];
        node_param_coords(HEX27, a, x);
        const int ir = x[0] > 0, is = x[1] > 0, it = x[2] > 0;
        const double R = Rf[ir], S = Sf[is], T = Tf[it];
        dN[0 + 3 * a] = (ir ? Q18 : -Q18) * S * T * (2.0 * R + S + T - 5.0);
        dN[1 + 3 * a] = (is ? Q18 : -Q18) * T * R * (2.0 * S + T + R - 5.0);
        dN[2 + 3 * a] = (it ? Q18 : -Q18) * R * S * (2.0 * T + R + S - 5.0);
      }
      const double v[12][3] = {{-0.5 * r * sm * tm, -0.25 * rrm * tm, -0.25 * rrm * sm}, {0.25 * ssm * tm, -0.5 * s * tm * rp, -0.25 * ssm * rp},
          {-0.5 * r * sp * tm, 0.25 * rrm * tm, -0.25 * rrm * sp}, {-0.25 * ssm * tm, -0.5 * s * tm * rm, -0.25 * ssm * rm},
          {-0.25 * sm * ttm, -0.25 * ttm * rm, -0.5 * t * rm * sm}, {0.25 * sm * ttm, -0.25 * ttm * rp, -0.5 * t * rp * sm},
          {0.25 * sp * ttm, 0.25 * ttm * rp, -0.5 * t * rp * sp}, {-0.25 * sp * ttm, 0.25 * ttm * rm, -0.5 * t * rm * sp},
          {-0.5 * r * sm * tp, -0.25 * rrm * tp, 0.25 * rrm * sm}, {0.25 * ssm * tp, -0.5 * s * tp * rp, 0.25 * ssm * rp},
          {-0.5 * r * sp * tp, 0.25 * rrm * tp, 0.25 * rrm * sp}, {-0.25 * ssm * tp, -0.5 * s * tp * rm, 0.25 * ssm * rm}};
      for (int a = 8; a < 20; ++a)
        for (int d = 0; d < 3; ++d) dN[d + 3 * a] = v[a - 8][d];
    }
    else if (ct == TET4)
    {
      const double v[4][3] = {{-1, -1, -1}, {1, 0, 0}, {0, 1, 0}, {0, 0, 1}};
      for (int a = 0; a < 4; ++a)
        for (int d = 0; d < 3; ++d) dN[d + 3 * a] = v[a][d];
    }
    else if (ct == WEDGE6)
    {
      const double Q12 = 1.0 / 2.0;
      const double p1 = Q12 * (1.0 - t), p2 = Q12 * (1.0 + t), t3 = 1.0 - r - s;
      const double v[6][3] = {{p1, 0, -Q12 * r}, {0, p1, -Q12 * s}, {-p1, -p1, -Q12 * t3}, {p2, 0, Q12 * r}, {0, p2, Q12 * s}, {-p2, -p2, Q12 * t3}};
      for (int a = 0; a < 6; ++a)
        for (int d = 0; d < 3; ++d) dN[d + 3 * a] = v[a][d];
    }
    else if (ct == PYRAMID5)
    {
      const double Q14 = 1.0 / 4.0;
      double rationdr = 1.0, rationds = 1.0, rationdt = 1.0;
      if (std::abs(t - 1.0) > 1.0e-14)
      {
        rationdr = s * t / (1.0 - t);
        rationds = r * t / (1.0 - t);
        rationdt = r * s / ((1.0 - t) * (1.0 - t));
      }
      const double v[5][3] = {{Q14 * (-1.0 * (1.0 - s) + rationdr), Q14 * (-1.0 * (1.0 - r) + rationds), Q14 * (rationdt - 1.0)},
          {Q14 * (1.0 * (1.0 - s) - rationdr), Q14 * (-1.0 * (1.0 + r) - rationds), Q14 * (-1.0 * rationdt - 1.0)},
          {Q14 * (1.0 * (1.0 + s) + rationdr), Q14 * (1.0 * (1.0 + r) + rationds), Q14 * (rationdt - 1.0)},
          {Q14 * (-1.0 * (1.0 + s) - rationdr), Q14 * (1.0 * (1.0 - r) - rationds), Q14 * (-1.0 * rationdt - 1.0)}, {0.0, 0.0, 1.0}};
      for (int a = 0; a < 5; ++a)
        for (int d = 0; d < 3; ++d) dN[d + 3 * a] = v[a][d];
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
    else if (ct == TET4)
    {
      xi[0][0] = xi[0][1] = xi[0][2] = 1.0 / 4.0;
      w[0] = 1.0 / 6.0;
    }
    else if (ct == WEDGE6)
    {
      const double xi3 = 1.0 / std::sqrt(3.0), Q23 = 2.0 / 3.0, Q16 = 1.0 / 6.0;
      for (int g = 0; g < 6; ++g)
      {
        xi[g][0] = g % 3 == 0 ? Q23 : Q16;
        xi[g][1] = g % 3 == 1 ? Q23 : Q16;
        xi[g][2] = g < 3 ? xi3 : -xi3;
        w[g] = Q16;
      }
    }
    else if (ct == PYRAMID5)
    {
      const double a = 0.26318405556971, b = 0.50661630334979, za = 0.54415184401122, zb = 0.12251482265544;
      const double wa = 0.10078588207983, wb = 0.23254745125351;
      const double p[8][3] = {{-a, -a, za}, {-b, -b, zb}, {-a, a, za}, {-b, b, zb}, {a, -a, za}, {b, -b, zb}, {a, a, za}, {b, b, zb}};
      for (int g = 0; g < 8; ++g)
      {
        for (int d = 0; d < 3; ++d) xi[g][d] = p[g][d];
        w[g] = (g & 1) ? wb : wa;
      }
    }
    else if (ct == HEX18)
    {
      // hex_18point (integration.cpp:112-130) = quad_9point (:6185-6233) x line_2point (:6946-6953), truncated constants
      const double xi3 = 0.7745966692415, w1 = 0.5555555555556, w2 = 0.8888888888889, xi2 = 0.5773502691896;
      for (int g = 0; g < 18; ++g)
      {
        double x[3];
        node_param_coords(HEX18, g, x);  // in-plane points ordered like the nodes
        xi[g][0] = x[0] == 0 ? 0.0 : (x[0] > 0 ? xi3 : -xi3);
        xi[g][1] = x[1] == 0 ? 0.0 : (x[1] > 0 ? xi3 : -xi3);
        xi[g][2] = g < 9 ? -xi2 : xi2;
        w[g] = 1.0 * ((x[0] == 0 ? w2 : w1) * (x[1] == 0 ? w2 : w1));
      }
    }
    else if (ct == HEX27 || ct == HEX20 || ct == NURBS27)  // nurbs27: hex_27point too (integration_select.hpp:89-92)
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

  // mass rule of a cell type where it differs from the stiffness rule (tet10 -> tet_11point, tet4 -> tet_4point);
  // returns the number of points, 0 = same rule as the stiffness integral
  inline int gauss_rule_mass(int ct, double (*xi)[3], double* w)
  {
    if (ct == TET10) return gauss_rule_mass_tet10(xi, w);
    if (ct == TET4)
    {
      gauss_rule(TET10, xi, w);
      return 4;
    }
    return 0;
  }

  // centroid in parameter space (used by F-bar), evaluate_parameter_coordinate_centroid
  // (src/solid_3D_ele/4C_solid_3D_ele_calc_lib.hpp:230-294): hex -> origin, tet -> (1/4,1/4,1/4), pyramid -> (0,0,1/4),
  // wedge -> (1/3,1/3,0)
  inline void centroid_param(int ct, double* xi)
  {
    const double c = (ct == TET10 || ct == TET4) ? 0.25 : 0.0;
    xi[0] = xi[1] = xi[2] = c;
    if (ct == PYRAMID5) xi[2] = 0.25;
    if (ct == WEDGE6) xi[0] = xi[1] = 1.0 / 3.0;
  }
}  // namespace fc_oracle
