// Surface cells of the Neumann loads (host + device): shape functions, first derivatives and Gauss rules of quad4 /
// quad9 / tri6 / quad8 / tri3, the faces of hex8 / hex27 / tet10 / hex20 / tet4 (wedge6 and pyramid5 have quad4 and
// tri3 faces).  Definitions follow the reference (paths relative to its tree):
//   shape functions  src/core/fem/src/general/utils/4C_fem_general_utils_fem_shapefunctions.hpp  quad4 :2003-2015,
//                    quad8 :2027-2064, quad9 :2065-2087, tri3 :2088-2097, tri6 :2098-2111 (derivatives :2148-, :2185-, :2233-,
//                    :2276-, :2288-)
//   Gauss rules      src/core/fem/src/general/utils/4C_fem_general_utils_integration.cpp  quad_4point :6151-6183,
//                    quad_9point :6185-6231 (quad8 and quad9), tri_3point :6341-6353, tri_6point :6374-6396 - with the
//                    reference's 13-digit constants
//   rule per cell    src/solid_3D_ele/4C_solid_3D_ele_surface.cpp:58-77 (get_optimal_gauss_rule)
// The shape functions are generated from the node positions (products of 1D Lagrange polynomials / barycentric
// coordinates) instead of the reference's per-node literals.
#pragma once
#include "fc_fem.cuh"

namespace fc
{
  enum
  {
    QUAD4 = 0,
    QUAD9 = 1,
    TRI6 = 2,
    QUAD8 = 3,
    TRI3 = 4,
    QUAD6 = 5  // side faces of hex18: nodes 0 (r = -1), 1 (r = +1), 2 (r = 0) at s = -1, then 3, 4, 5 at s = +1
  };

  template <int FT>
  struct FaceCell
  {
    static constexpr int nen = FT == QUAD4 ? 4 : (FT == QUAD9 ? 9 : (FT == TRI6 || FT == QUAD6 ? 6 : (FT == QUAD8 ? 8 : 3)));
    static constexpr int ngp = FT == QUAD4 ? 4 : (FT == QUAD9 || FT == QUAD8 ? 9 : (FT == TRI6 || FT == QUAD6 ? 6 : 3));
  };
  FC_HD int face_nen(int ft) { return ft == QUAD4 ? 4 : (ft == QUAD9 ? 9 : (ft == TRI6 || ft == QUAD6 ? 6 : (ft == QUAD8 ? 8 : 3))); }
  FC_HD int face_ngp(int ft) { return ft == QUAD4 ? 4 : (ft == QUAD9 || ft == QUAD8 ? 9 : (ft == TRI6 || ft == QUAD6 ? 6 : 3)); }
  constexpr int face_max_nen = 9, face_max_ngp = 9;

  // in-face position of quad node a in direction d as an integer in {-1, 0, 1}: corners 0-3 counter-clockwise from
  // (-1,-1), edge mids 4-7 starting on the edge 0-1, centre 8
  FC_HD int quad_node_pos(int a, int d)
  {
    const signed char r[9] = {-1, 1, 1, -1, 0, 1, 0, -1, 0}, s[9] = {-1, -1, 1, 1, -1, 0, 1, 0, 0};
    return d == 0 ? r[a] : s[a];
  }

  // point g of the surface rule of face type ft
  FC_HD void face_gauss_point(int ft, int g, double* rs, double* w)
  {
    if (ft == QUAD4)
    {
      const double p = 0.5773502691896;
      rs[0] = quad_node_pos(g, 0) * p;
      rs[1] = quad_node_pos(g, 1) * p;
      *w = 1.0;
    }
    else if (ft == TRI3)
    {
      rs[0] = g == 1 ? 2.0 / 3.0 : 1.0 / 6.0;
      rs[1] = g == 2 ? 2.0 / 3.0 : 1.0 / 6.0;
      *w = 1.0 / 6.0;
    }
    else if (ft == QUAD9 || ft == QUAD8)
    {
      // the points sit where the nodes of a quad9 sit (scaled), in the same order
      const double p = 0.7745966692415, w1 = 0.5555555555556, w2 = 0.8888888888889;
      const int a = quad_node_pos(g, 0), b = quad_node_pos(g, 1);
      rs[0] = a * p;
      rs[1] = b * p;
      // the reference writes weight(xi_1) * weight(xi_2)
      *w = (a == 0 ? w2 : w1) * (b == 0 ? w2 : w1);
    }
    else if (ft == QUAD6)
    {
      // quad_6point (4C_fem_general_utils_integration.cpp:6235-6261): three points in r, two in s, in the order of the nodes
      const double p = 0.7745966692415, q = 0.5773502691896;
      const int k = g % 3;
      rs[0] = k == 0 ? -p : (k == 1 ? p : 0.0);
      rs[1] = g < 3 ? -q : q;
      *w = k == 2 ? 0.8888888888889 : 0.5555555555556;
    }
    else
    {
      const double a = 0.0915762135098, b = 0.8168475729805, c = 0.4459484909160, d = 0.1081030181681;
      const double pr[6] = {a, b, a, c, c, d}, ps[6] = {a, a, b, d, c, c};
      rs[0] = pr[g];
      rs[1] = ps[g];
      *w = g < 3 ? 0.0549758718277 : 0.1116907948390;
    }
  }

  // N[a], dN/dr[a], dN/ds[a] at (r, s)
  FC_HD void face_shape(int ft, const double* rs, double* N, double* dNr, double* dNs)
  {
    const double r = rs[0], s = rs[1];
    if (ft == QUAD4)
    {
      for (int a = 0; a < 4; ++a)
      {
        const double ra = quad_node_pos(a, 0), sa = quad_node_pos(a, 1);
        N[a] = 0.25 * (1.0 + ra * r) * (1.0 + sa * s);
        dNr[a] = 0.25 * ra * (1.0 + sa * s);
        dNs[a] = 0.25 * (1.0 + ra * r) * sa;
      }
    }
    else if (ft == QUAD8)
    {
      // serendipity: corner = bilinear - half of the two neighbouring mid-edge functions
      const double r2 = 1.0 - r * r, s2 = 1.0 - s * s;
      for (int a = 0; a < 8; ++a)
      {
        const double ra = quad_node_pos(a, 0), sa = quad_node_pos(a, 1);
        const double R = 1.0 + ra * r, S = 1.0 + sa * s;
        if (a < 4)
        {
          N[a] = 0.25 * (R * S - (r2 * S + s2 * R));
          dNr[a] = 0.25 * (ra * S - (-2.0 * r * S + s2 * ra));
          dNs[a] = 0.25 * (R * sa - (r2 * sa - 2.0 * s * R));
        }
        else if (ra == 0.0)
        {
          N[a] = 0.5 * r2 * S;
          dNr[a] = -r * S;
          dNs[a] = 0.5 * r2 * sa;
        }
        else
        {
          N[a] = 0.5 * s2 * R;
          dNr[a] = 0.5 * s2 * ra;
          dNs[a] = -s * R;
        }
      }
    }
    else if (ft == TRI3)
    {
      N[0] = 1.0 - r - s;
      N[1] = r;
      N[2] = s;
      dNr[0] = -1.0;
      dNr[1] = 1.0;
      dNr[2] = 0.0;
      dNs[0] = -1.0;
      dNs[1] = 0.0;
      dNs[2] = 1.0;
    }
    else if (ft == QUAD9)
    {
      const double Lr[3] = {0.5 * r * (r - 1.0), 1.0 - r * r, 0.5 * r * (r + 1.0)}, dLr[3] = {r - 0.5, -2.0 * r, r + 0.5};
      const double Ls[3] = {0.5 * s * (s - 1.0), 1.0 - s * s, 0.5 * s * (s + 1.0)}, dLs[3] = {s - 0.5, -2.0 * s, s + 0.5};
      for (int a = 0; a < 9; ++a)
      {
        const int ir = quad_node_pos(a, 0) + 1, is = quad_node_pos(a, 1) + 1;
        N[a] = Lr[ir] * Ls[is];
        dNr[a] = dLr[ir] * Ls[is];
        dNs[a] = Lr[ir] * dLs[is];
      }
    }
    else if (ft == QUAD6)
    {
      // 4C_fem_general_utils_fem_shapefunctions.hpp:2016-2026, 2168-2184
      const double Lr[3] = {0.5 * r * (r - 1.0), 0.5 * r * (r + 1.0), 1.0 - r * r}, dLr[3] = {r - 0.5, r + 0.5, -2.0 * r};
      for (int a = 0; a < 6; ++a)
      {
        const double Ls = a < 3 ? 0.5 * (1.0 - s) : 0.5 * (1.0 + s);
        N[a] = Lr[a % 3] * Ls;
        dNr[a] = dLr[a % 3] * Ls;
        dNs[a] = Lr[a % 3] * (a < 3 ? -0.5 : 0.5);
      }
    }
    else
    {
      // barycentric coordinates (t1, t2, t3) = (1 - r - s, r, s); corners 0-2, mids 3: (0,1), 4: (1,2), 5: (2,0)
      const double t[3] = {1.0 - r - s, r, s};
      const double dtr[3] = {-1.0, 1.0, 0.0}, dts[3] = {-1.0, 0.0, 1.0};
      for (int a = 0; a < 3; ++a)
      {
        N[a] = t[a] * (2.0 * t[a] - 1.0);
        dNr[a] = (4.0 * t[a] - 1.0) * dtr[a];
        dNs[a] = (4.0 * t[a] - 1.0) * dts[a];
      }
      for (int m = 0; m < 3; ++m)
      {
        const int i = m, j = (m + 1) % 3;
        N[3 + m] = 4.0 * t[i] * t[j];
        dNr[3 + m] = 4.0 * (t[i] * dtr[j] + t[j] * dtr[i]);
        dNs[3 + m] = 4.0 * (t[i] * dts[j] + t[j] * dts[i]);
      }
    }
  }
}  // namespace fc
