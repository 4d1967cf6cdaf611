// ORACLE (test infrastructure): CPU restatement of the reference's Neumann loads of TYPE Live.
//   surface  Discret::Elements::SolidSurface::evaluate_neumann, case neum_live
//            src/solid_3D_ele/4C_solid_3D_ele_surface_evaluate.cpp:225-330
//   volume   Discret::Elements::evaluate_neumann
//            src/solid_3D_ele/4C_solid_3D_ele_neumann_evaluator.cpp:43-115
//   2D shape functions  src/core/fem/src/general/utils/4C_fem_general_utils_fem_shapefunctions.hpp:2003-2015 (quad4),
//                       :2027-2064 (quad8), :2065-2087 (quad9), :2088-2097 (tri3), :2098-2111 (tri6) and their first
//                       derivatives (:2148-, :2185-2232, :2233-, :2276-2287, :2288-)
//   2D Gauss rules      src/core/fem/src/general/utils/4C_fem_general_utils_integration.cpp:6151-6183 (quad_4point),
//                       :6185-6231 (quad_9point, also the rule of quad8), :6341-6353 (tri_3point), :6374-6396 (tri_6point),
//                       truncated constants as written there
//   rule per cell       src/solid_3D_ele/4C_solid_3D_ele_surface.cpp:58-77
// Parity: pinned by the reference's RESULT DESCRIPTION of tests/input_files/solid_ele_hex8_Standard_eas_none_volume_neumann.dat
// (surface + volume loads with FUNCT (x+1)(y+2)(z+3) and t; tests/golden) and by the surface loads of the other goldens.
#pragma once
#include <cmath>

#include "fc_oracle_fem.hpp"

namespace fc_oracle
{
  enum
  {
    QUAD4 = 0,
    QUAD9 = 1,
    TRI6 = 2,
    QUAD8 = 3,
    TRI3 = 4,
    QUAD6 = 5,
    NURBS9 = 6  // faces of nurbs27: knots [2][6] and weights [9] of the face in nurbs_context() (fc_oracle_nurbs.hpp);
                // 4C_solid_3D_ele_surface_evaluate.cpp:190-256, quad_9point like quad9 (4C_solid_3D_ele_surface.cpp:68-71)
  };
  inline int face_num_nodes(int ft)
  {
    static const int n[7] = {4, 9, 6, 8, 3, 6, 9};
    return n[ft];
  }
  inline int face_num_gp(int ft)
  {
    static const int n[7] = {4, 9, 6, 9, 3, 6, 9};
    return n[ft];
  }

  // shape_function_2d / shape_function_2d_deriv1
  inline void shape_function_2d(int ft, double r, double s, double* funct, double* deriv_r, double* deriv_s)
  {
    if (ft == NURBS9)
    {
      const double uv[2] = {r, s};
      double d[18];
      nurbs_funct_deriv<2>(uv, nurbs_context().knots, nurbs_context().weights, funct, d);
      for (int a = 0; a < 9; ++a)
      {
        deriv_r[a] = d[2 * a];
        deriv_s[a] = d[2 * a + 1];
      }
    }
    else if (ft == QUAD4)
    {
      const double rp = 1.0 + r, rm = 1.0 - r, sp = 1.0 + s, sm = 1.0 - s;
      funct[0] = 0.25 * rm * sm;
      funct[1] = 0.25 * rp * sm;
      funct[2] = 0.25 * rp * sp;
      funct[3] = 0.25 * rm * sp;
      deriv_r[0] = -0.25 * sm;
      deriv_s[0] = -0.25 * rm;
      deriv_r[1] = 0.25 * sm;
      deriv_s[1] = -0.25 * rp;
      deriv_r[2] = 0.25 * sp;
      deriv_s[2] = 0.25 * rp;
      deriv_r[3] = -0.25 * sp;
      deriv_s[3] = 0.25 * rm;
    }
    else if (ft == QUAD8)
    {
      const double rp = 1.0 + r, rm = 1.0 - r, sp = 1.0 + s, sm = 1.0 - s;
      const double r2 = 1.0 - r * r, s2 = 1.0 - s * s;
      funct[0] = 0.25 * (rm * sm - (r2 * sm + s2 * rm));
      funct[1] = 0.25 * (rp * sm - (r2 * sm + s2 * rp));
      funct[2] = 0.25 * (rp * sp - (s2 * rp + r2 * sp));
      funct[3] = 0.25 * (rm * sp - (r2 * sp + s2 * rm));
      funct[4] = 0.5 * r2 * sm;
      funct[5] = 0.5 * s2 * rp;
      funct[6] = 0.5 * r2 * sp;
      funct[7] = 0.5 * s2 * rm;
      deriv_r[0] = 0.25 * sm * (2. * r + s);
      deriv_s[0] = 0.25 * rm * (r + 2. * s);
      deriv_r[1] = 0.25 * sm * (2. * r - s);
      deriv_s[1] = 0.25 * rp * (2. * s - r);
      deriv_r[2] = 0.25 * sp * (2. * r + s);
      deriv_s[2] = 0.25 * rp * (r + 2. * s);
      deriv_r[3] = 0.25 * sp * (2. * r - s);
      deriv_s[3] = 0.25 * rm * (2. * s - r);
      deriv_r[4] = -sm * r;
      deriv_s[4] = -0.5 * rm * rp;
      deriv_r[5] = 0.5 * sm * sp;
      deriv_s[5] = -rp * s;
      deriv_r[6] = -sp * r;
      deriv_s[6] = 0.5 * rm * rp;
      deriv_r[7] = -0.5 * sm * sp;
      deriv_s[7] = -rm * s;
    }
    else if (ft == QUAD6)
    {
      // side face of hex18 (4C_fem_general_utils_fem_shapefunctions.hpp:2016-2026, 2168-2184): quadratic in r, linear in s
      funct[0] = 0.25 * r * (r - 1.) * (1. - s);
      funct[1] = 0.25 * r * (r + 1.) * (1. - s);
      funct[2] = -0.5 * (r - 1.) * (r + 1.) * (1. - s);
      funct[3] = 0.25 * r * (r - 1.) * (1. + s);
      funct[4] = 0.25 * r * (r + 1.) * (1. + s);
      funct[5] = -0.5 * (r - 1.) * (r + 1.) * (1. + s);
      deriv_r[0] = 0.25 * (2. * r - 1.) * (1. - s);
      deriv_r[1] = 0.25 * (2. * r + 1.) * (1. - s);
      deriv_r[2] = -r * (1. - s);
      deriv_r[3] = 0.25 * (2. * r - 1.) * (1. + s);
      deriv_r[4] = 0.25 * (2. * r + 1.) * (1. + s);
      deriv_r[5] = -r * (1. + s);
      deriv_s[0] = -0.25 * r * (r - 1.);
      deriv_s[1] = -0.25 * r * (r + 1.);
      deriv_s[2] = +0.5 * (r - 1.) * (r + 1.);
      deriv_s[3] = 0.25 * r * (r - 1.);
      deriv_s[4] = 0.25 * r * (r + 1.);
      deriv_s[5] = -0.5 * (r - 1.) * (r + 1.);
    }
    else if (ft == TRI3)
    {
      funct[0] = 1.0 - r - s;
      funct[1] = r;
      funct[2] = s;
      deriv_r[0] = -1.0;
      deriv_r[1] = 1.0;
      deriv_r[2] = 0.0;
      deriv_s[0] = -1.0;
      deriv_s[1] = 0.0;
      deriv_s[2] = 1.0;
    }
    else if (ft == QUAD9)
    {
      const double rp = 1.0 + r, rm = 1.0 - r, sp = 1.0 + s, sm = 1.0 - s;
      const double r2 = 1.0 - r * r, s2 = 1.0 - s * s, rh = 0.5 * r, sh = 0.5 * s, rs = rh * sh;
      const double rhp = r + 0.5, rhm = r - 0.5, shp = s + 0.5, shm = s - 0.5;
      funct[0] = rs * rm * sm;
      funct[1] = -rs * rp * sm;
      funct[2] = rs * rp * sp;
      funct[3] = -rs * rm * sp;
      funct[4] = -sh * sm * r2;
      funct[5] = rh * rp * s2;
      funct[6] = sh * sp * r2;
      funct[7] = -rh * rm * s2;
      funct[8] = r2 * s2;
      deriv_r[0] = -rhm * sh * sm;
      deriv_s[0] = -shm * rh * rm;
      deriv_r[1] = -rhp * sh * sm;
      deriv_s[1] = shm * rh * rp;
      deriv_r[2] = rhp * sh * sp;
      deriv_s[2] = shp * rh * rp;
      deriv_r[3] = rhm * sh * sp;
      deriv_s[3] = -shp * rh * rm;
      deriv_r[4] = 2.0 * r * sh * sm;
      deriv_s[4] = shm * r2;
      deriv_r[5] = rhp * s2;
      deriv_s[5] = -2.0 * s * rh * rp;
      deriv_r[6] = -2.0 * r * sh * sp;
      deriv_s[6] = shp * r2;
      deriv_r[7] = rhm * s2;
      deriv_s[7] = 2.0 * s * rh * rm;
      deriv_r[8] = -2.0 * r * s2;
      deriv_s[8] = -2.0 * s * r2;
    }
    else
    {
      const double t1 = 1.0 - r - s, t2 = r, t3 = s;
      funct[0] = t1 * (2.0 * t1 - 1.0);
      funct[1] = t2 * (2.0 * t2 - 1.0);
      funct[2] = t3 * (2.0 * t3 - 1.0);
      funct[3] = 4.0 * t2 * t1;
      funct[4] = 4.0 * t2 * t3;
      funct[5] = 4.0 * t3 * t1;
      deriv_r[0] = -3.0 + 4.0 * (r + s);
      deriv_s[0] = -3.0 + 4.0 * (r + s);
      deriv_r[1] = 4.0 * r - 1.0;
      deriv_s[1] = 0.0;
      deriv_r[2] = 0.0;
      deriv_s[2] = 4.0 * s - 1.0;
      deriv_r[3] = 4.0 * (1.0 - 2.0 * r - s);
      deriv_s[3] = -4.0 * r;
      deriv_r[4] = 4.0 * s;
      deriv_s[4] = 4.0 * r;
      deriv_r[5] = -4.0 * s;
      deriv_s[5] = 4.0 * (1.0 - r - 2.0 * s);
    }
  }

  // IntegrationPoints2D(gaussrule_)
  inline void gauss_rule_2d(int ft, double (*qxg)[2], double* qwgt)
  {
    if (ft == QUAD4)
    {
      const double g = 0.5773502691896;
      const double p[4][2] = {{-g, -g}, {g, -g}, {g, g}, {-g, g}};
      for (int i = 0; i < 4; ++i)
      {
        qxg[i][0] = p[i][0];
        qxg[i][1] = p[i][1];
        qwgt[i] = 1.0;
      }
    }
    else if (ft == TRI3)
    {
      const double p[3][2] = {{1.0 / 6.0, 1.0 / 6.0}, {2.0 / 3.0, 1.0 / 6.0}, {1.0 / 6.0, 2.0 / 3.0}};
      for (int i = 0; i < 3; ++i)
      {
        qxg[i][0] = p[i][0];
        qxg[i][1] = p[i][1];
        qwgt[i] = 1.0 / 6.0;
      }
    }
    else if (ft == QUAD6)
    {
      // quad_6point, 4C_fem_general_utils_integration.cpp:6235-6261
      const double g = 0.7745966692415, h = 0.5773502691896, a = 0.5555555555556, b = 0.8888888888889;
      const double p[6][2] = {{-g, -h}, {g, -h}, {0.0, -h}, {-g, h}, {g, h}, {0.0, h}};
      for (int i = 0; i < 6; ++i)
      {
        qxg[i][0] = p[i][0];
        qxg[i][1] = p[i][1];
        qwgt[i] = (i % 3 == 2) ? b : a;
      }
    }
    else if (ft == QUAD9 || ft == QUAD8 || ft == NURBS9)
    {
      const double g = 0.7745966692415, a = 0.5555555555556, b = 0.8888888888889;
      const double p[9][2] = {{-g, -g}, {g, -g}, {g, g}, {-g, g}, {0.0, -g}, {g, 0.0}, {0.0, g}, {-g, 0.0}, {0.0, 0.0}};
      const double w[9] = {a * a, a * a, a * a, a * a, b * a, a * b, b * a, a * b, b * b};
      for (int i = 0; i < 9; ++i)
      {
        qxg[i][0] = p[i][0];
        qxg[i][1] = p[i][1];
        qwgt[i] = w[i];
      }
    }
    else
    {
      const double p[6][2] = {{0.0915762135098, 0.0915762135098}, {0.8168475729805, 0.0915762135098}, {0.0915762135098, 0.8168475729805},
          {0.4459484909160, 0.1081030181681}, {0.4459484909160, 0.4459484909160}, {0.1081030181681, 0.4459484909160}};
      for (int i = 0; i < 6; ++i)
      {
        qxg[i][0] = p[i][0];
        qxg[i][1] = p[i][1];
        qwgt[i] = i < 3 ? 0.0549758718277 : 0.1116907948390;
      }
    }
  }

  // x [numnode][3] reference coordinates of the face nodes; functfac [ngp][3] or nullptr (= 1); elevec1 [numnode][3] +=
  // gp_coord [ngp][3] (optional output): reference position of the Gauss points (multiply_tn(gp_coord, funct, x))
  inline void surface_neumann_live(int ft, const double* x, const int* onoff, const double* val, const double* functfac, double* elevec1,
      double* gp_coord)
  {
    const int numnode = face_num_nodes(ft), nquad = face_num_gp(ft);
    double qxg[9][2], qwgt[9];
    gauss_rule_2d(ft, qxg, qwgt);
    for (int gp = 0; gp < nquad; ++gp)
    {
      double funct[9], deriv[2][9];
      shape_function_2d(ft, qxg[gp][0], qxg[gp][1], funct, deriv[0], deriv[1]);
      if (gp_coord)
        for (int i = 0; i < 3; ++i)
        {
          double c = 0.0;
          for (int n = 0; n < numnode; ++n) c += funct[n] * x[3 * n + i];
          gp_coord[3 * gp + i] = c;
        }
      if (!elevec1) continue;
      double dxyzdrs[2][3] = {{0, 0, 0}, {0, 0, 0}};
      for (int a = 0; a < 2; ++a)
        for (int i = 0; i < 3; ++i)
          for (int n = 0; n < numnode; ++n) dxyzdrs[a][i] += deriv[a][n] * x[3 * n + i];
      double metrictensor[2][2];
      for (int a = 0; a < 2; ++a)
        for (int b = 0; b < 2; ++b)
        {
          metrictensor[a][b] = 0.0;
          for (int i = 0; i < 3; ++i) metrictensor[a][b] += dxyzdrs[a][i] * dxyzdrs[b][i];
        }
      const double detA = std::sqrt(metrictensor[0][0] * metrictensor[1][1] - metrictensor[0][1] * metrictensor[1][0]);
      for (int dof = 0; dof < 3; ++dof)
        if (onoff[dof])
        {
          const double ff = functfac ? functfac[3 * gp + dof] : 1.0;
          const double fac = qwgt[gp] * detA * val[dof] * ff;
          for (int node = 0; node < numnode; ++node) elevec1[node * 3 + dof] += funct[node] * fac;
        }
    }
  }

  // TYPE pseudo_orthopressure / orthopressure (4C_solid_3D_ele_surface_evaluate.cpp:318-358): xc [numnode][3] = the configuration the
  // load type asks for (:139-176); ortho_value = val[0]; normal of surface_integration (:382-394), its length the area element
  inline void surface_neumann_orthopressure(int ft, const double* xc, double ortho_value, const double* functfac, double* elevec1)
  {
    const int numnode = face_num_nodes(ft), nquad = face_num_gp(ft);
    double qxg[9][2], qwgt[9];
    gauss_rule_2d(ft, qxg, qwgt);
    for (int gp = 0; gp < nquad; ++gp)
    {
      double funct[9], deriv[2][9];
      shape_function_2d(ft, qxg[gp][0], qxg[gp][1], funct, deriv[0], deriv[1]);
      double dxyzdrs[2][3] = {{0, 0, 0}, {0, 0, 0}};
      for (int a = 0; a < 2; ++a)
        for (int i = 0; i < 3; ++i)
          for (int n = 0; n < numnode; ++n) dxyzdrs[a][i] += deriv[a][n] * xc[3 * n + i];
      const double normal[3] = {dxyzdrs[0][1] * dxyzdrs[1][2] - dxyzdrs[0][2] * dxyzdrs[1][1],
          dxyzdrs[0][2] * dxyzdrs[1][0] - dxyzdrs[0][0] * dxyzdrs[1][2], dxyzdrs[0][0] * dxyzdrs[1][1] - dxyzdrs[0][1] * dxyzdrs[1][0]};
      const double fac = qwgt[gp] * (functfac ? functfac[3 * gp] : 1.0) * ortho_value;  // normalfac = 1 off NURBS
      for (int node = 0; node < numnode; ++node)
        for (int dim = 0; dim < 3; ++dim) elevec1[node * 3 + dim] += funct[node] * normal[dim] * fac;
    }
  }

  // Load linearization of the orthopressure load (LOADLIN yes; surface_evaluate.cpp:358-376, analytical_d_surface_integration :502-549):
  // elemat1 [3 numnode][3 numnode] row-major, (node*3 + dim, dof) -= funct[node] * Dnormal(dim, dof) * fac
  inline void surface_neumann_orthopressure_linearization(int ft, const double* xc, double ortho_value, const double* functfac, double* elemat1)
  {
    const int numnode = face_num_nodes(ft), nquad = face_num_gp(ft), numdof = 3 * numnode;
    double qxg[9][2], qwgt[9];
    gauss_rule_2d(ft, qxg, qwgt);
    for (int gp = 0; gp < nquad; ++gp)
    {
      double funct[9], deriv[2][9];
      shape_function_2d(ft, qxg[gp][0], qxg[gp][1], funct, deriv[0], deriv[1]);
      double g1[3] = {0, 0, 0}, g2[3] = {0, 0, 0};
      for (int i = 0; i < 3; ++i)
        for (int n = 0; n < numnode; ++n)
        {
          g1[i] += deriv[0][n] * xc[3 * n + i];
          g2[i] += deriv[1][n] * xc[3 * n + i];
        }
      const double fac = qwgt[gp] * (functfac ? functfac[3 * gp] : 1.0) * ortho_value;
      for (int dof = 0; dof < numdof; ++dof)
      {
        const int nb = dof / 3, k = dof % 3;
        double dg1[3] = {0, 0, 0}, dg2[3] = {0, 0, 0};
        dg1[k] = deriv[0][nb];
        dg2[k] = deriv[1][nb];
        const double dn[3] = {dg1[1] * g2[2] + g1[1] * dg2[2] - dg1[2] * g2[1] - g1[2] * dg2[1],
            dg1[2] * g2[0] + g1[2] * dg2[0] - dg1[0] * g2[2] - g1[0] * dg2[2], dg1[0] * g2[1] + g1[0] * dg2[1] - dg1[1] * g2[0] - g1[1] * dg2[0]};
        for (int node = 0; node < numnode; ++node)
          for (int dim = 0; dim < 3; ++dim) elemat1[(size_t)(node * 3 + dim) * numdof + dof] -= funct[node] * dn[dim] * fac;
      }
    }
  }

  // X [nen][3]; element_force_vector [nen][3] +=; gauss rule = get_gauss_rule_stiffness_matrix<celltype>
  inline void volume_neumann_live(int ct, const double* X, const int* onoff, const double* value, const double* functfac,
      double* element_force_vector, double* gp_coord)
  {
    const int numnod = num_nodes(ct), ngp = num_gp(ct);
    double xi[27][3], w[27];
    gauss_rule(ct, xi, w);
    for (int gp = 0; gp < ngp; ++gp)
    {
      double N[27], dN[81];
      shape_function(ct, xi[gp], N);
      shape_function_deriv1(ct, xi[gp], dN);
      // jacobian_mapping: J(d, i) = sum_a dN(d, a) X(a, i); integration_factor = w det J
      double J[3][3] = {{0, 0, 0}, {0, 0, 0}, {0, 0, 0}};
      for (int d = 0; d < 3; ++d)
        for (int i = 0; i < 3; ++i)
          for (int a = 0; a < numnod; ++a) J[d][i] += dN[3 * a + d] * X[3 * a + i];
      const double det = J[0][0] * (J[1][1] * J[2][2] - J[1][2] * J[2][1]) - J[0][1] * (J[1][0] * J[2][2] - J[1][2] * J[2][0]) +
                         J[0][2] * (J[1][0] * J[2][1] - J[1][1] * J[2][0]);
      const double integration_factor = w[gp] * det;
      if (gp_coord)
        for (int i = 0; i < 3; ++i)
        {
          double c = 0.0;
          for (int a = 0; a < numnod; ++a) c += X[3 * a + i] * N[a];
          gp_coord[3 * gp + i] = c;
        }
      if (!element_force_vector) continue;
      for (int i = 0; i < 3; ++i)
        if (onoff[i])
        {
          const double function_scale_factor = functfac ? functfac[3 * gp + i] : 1.0;
          const double value_times_integration_factor = value[i] * function_scale_factor * integration_factor;
          for (int nodeid = 0; nodeid < numnod; ++nodeid) element_force_vector[nodeid * 3 + i] += N[nodeid] * value_times_integration_factor;
        }
    }
  }
}  // namespace fc_oracle
