// Neumann loads of TYPE Live: dead loads integrated over the reference configuration.
//   surface  src/solid_3D_ele/4C_solid_3D_ele_surface_evaluate.cpp:225-330 (neum_live):
//            f[a][dof] += N_a w sqrt(det(metric)) val[dof] funct[dof](X_gp, t)
//   volume   src/solid_3D_ele/4C_solid_3D_ele_neumann_evaluator.cpp:43-115 (evaluate_neumann):
//            f[a][dof] += N_a w detJ val[dof] funct[dof](X_gp, t), stiffness Gauss rule of the cell
// The function factor funct[dof](X_gp, t) is the caller's: gp_factor[item][gp][3] (nullptr = 1), evaluated at the
// reference coordinates that k_neumann_*_gp_coords hand out.  Owned rows only, like the element path.
// Not on the per-iteration path (loads change per time step): one thread per face / element, atomics into f_ext.
#pragma once
#include "fc_fem_surface.cuh"
#include "fc_kernels_common.cuh"
#include "fc_nurbs.cuh"

namespace fc
{
  struct NeumannVal
  {
    double val[3];
    int onoff[3];
  };

  // One instance per face type: every loop bound is a compile-time constant.  (A single kernel with the face type as a
  // run-time argument - bounds read from a table - was miscompiled by nvcc 12.9 -O3 for tri6: illegal address in the
  // accumulation loop, clean with -G.)
  // ORTHO: TYPE pseudo_orthopressure / orthopressure (4C_solid_3D_ele_surface_evaluate.cpp:84-96, 139-176, 318-358): the pressure
  // V.val[0] acts along the normal of the configuration X + u (u = the last converged displacements for pseudo_orthopressure, the
  // current iterate for orthopressure, none while MULF prestressing is active); the normal is the cross product of the two
  // tangents, its length the area element (surface_integration, :382-394):  f[a][dim] += N_a normal[dim] w functfac val[0]
  template <int FT, bool ORTHO = false>
  static __global__ void k_neumann_surface(const double* __restrict__ X, const int* __restrict__ face_nodes, int n_faces, NeumannVal V,
      const double* __restrict__ gp_factor, int n_row_nodes, double* __restrict__ f, double* __restrict__ gp_coords,
      const double* __restrict__ u = nullptr)
  {
    const int face = blockIdx.x * blockDim.x + threadIdx.x;
    if (face >= n_faces) return;
    constexpr int ft = FT, nen = FaceCell<FT>::nen, ngp = FaceCell<FT>::ngp;
    double P[nen][3];
    int nodes[nen];
    for (int a = 0; a < nen; ++a)
    {
      nodes[a] = face_nodes[(size_t)face * nen + a];
      for (int i = 0; i < 3; ++i) P[a][i] = X[3 * (size_t)nodes[a] + i] + (ORTHO && u ? u[3 * (size_t)nodes[a] + i] : 0.0);
    }
    double fa[nen][3];
    for (int a = 0; a < nen; ++a) fa[a][0] = fa[a][1] = fa[a][2] = 0.0;
    for (int g = 0; g < ngp; ++g)
    {
      double rs[2], w, N[nen], dr[nen], ds[nen];
      face_gauss_point(ft, g, rs, &w);
      face_shape(ft, rs, N, dr, ds);
      if (gp_coords)
      {
        for (int i = 0; i < 3; ++i)
        {
          double x = 0.0;
          for (int a = 0; a < nen; ++a) x += N[a] * P[a][i];
          gp_coords[((size_t)face * ngp + g) * 3 + i] = x;
        }
        continue;
      }
      double g1[3] = {0, 0, 0}, g2[3] = {0, 0, 0};
      for (int a = 0; a < nen; ++a)
        for (int i = 0; i < 3; ++i)
        {
          g1[i] += dr[a] * P[a][i];
          g2[i] += ds[a] * P[a][i];
        }
      if (ORTHO)
      {
        const double normal[3] = {g1[1] * g2[2] - g1[2] * g2[1], g1[2] * g2[0] - g1[0] * g2[2], g1[0] * g2[1] - g1[1] * g2[0]};
        const double functfac = gp_factor ? gp_factor[((size_t)face * ngp + g) * 3] : 1.0;
        const double fac = w * functfac * V.val[0];
        for (int a = 0; a < nen; ++a)
          for (int dim = 0; dim < 3; ++dim) fa[a][dim] += N[a] * normal[dim] * fac;
        continue;
      }
      const double m11 = g1[0] * g1[0] + g1[1] * g1[1] + g1[2] * g1[2];
      const double m22 = g2[0] * g2[0] + g2[1] * g2[1] + g2[2] * g2[2];
      const double m12 = g1[0] * g2[0] + g1[1] * g2[1] + g1[2] * g2[2];
      const double detA = sqrt(m11 * m22 - m12 * m12);
      for (int dof = 0; dof < 3; ++dof)
      {
        if (!V.onoff[dof]) continue;
        const double functfac = gp_factor ? gp_factor[((size_t)face * ngp + g) * 3 + dof] : 1.0;
        const double fac = w * detA * V.val[dof] * functfac;
        for (int a = 0; a < nen; ++a) fa[a][dof] += N[a] * fac;
      }
    }
    if (gp_coords) return;
    for (int a = 0; a < nen; ++a)
      if (nodes[a] < n_row_nodes)
        for (int dof = 0; dof < 3; ++dof)
          if (ORTHO || V.onoff[dof]) red_add(f + 3 * (size_t)nodes[a] + dof, fa[a][dof]);
  }

  // Load linearization of the orthopressure load (LOADLIN yes; 4C_solid_3D_ele_surface_evaluate.cpp:358-376 with
  // analytical_d_surface_integration :502-549): K[(a, dim)][(b, k)] -= N_a d normal[dim] / d x_(b,k) w functfac val[0], the
  // derivative of the cross product g1 x g2 with g1 = sum_b dr_b x_b, g2 = sum_b ds_b x_b.  One thread per face; every Gauss point
  // adds its entries into the owned, non-Dirichlet rows of the CSR values (position of node b in the row of node a by bisection
  // in the node graph).  Unsymmetric, like the reference's.
  template <int FT>
  static __global__ void k_orthopressure_linearization(const double* __restrict__ X, const double* __restrict__ u, const int* __restrict__ face_nodes,
      int n_faces, double pressure, const double* __restrict__ gp_factor, int n_row_nodes, const long long* __restrict__ nodeptr,
      const int* __restrict__ nodeadj, const int64_t* __restrict__ node_rowbase, const int* __restrict__ node_rowlen,
      const unsigned char* __restrict__ dbc, double* __restrict__ K, int* __restrict__ missing)
  {
    const int face = blockIdx.x * blockDim.x + threadIdx.x;
    if (face >= n_faces) return;
    constexpr int ft = FT, nen = FaceCell<FT>::nen, ngp = FaceCell<FT>::ngp;
    double P[nen][3];
    int nodes[nen], pos[nen][nen];
    for (int a = 0; a < nen; ++a)
    {
      nodes[a] = face_nodes[(size_t)face * nen + a];
      for (int i = 0; i < 3; ++i) P[a][i] = X[3 * (size_t)nodes[a] + i] + (u ? u[3 * (size_t)nodes[a] + i] : 0.0);
    }
    for (int a = 0; a < nen; ++a)
      for (int b = 0; b < nen; ++b)
      {
        pos[a][b] = -1;
        if (nodes[a] >= n_row_nodes) continue;
        long long lo = nodeptr[nodes[a]], hi = nodeptr[nodes[a] + 1] - 1;
        const long long base = lo;
        while (lo <= hi)
        {
          const long long mid = (lo + hi) >> 1;
          const int v = nodeadj[mid];
          if (v == nodes[b])
          {
            pos[a][b] = (int)(mid - base);
            break;
          }
          if (v < nodes[b])
            lo = mid + 1;
          else
            hi = mid - 1;
        }
        if (pos[a][b] < 0) atomicAdd(missing, 1);
      }
    for (int g = 0; g < ngp; ++g)
    {
      double rs[2], w, N[nen], dr[nen], ds[nen];
      face_gauss_point(ft, g, rs, &w);
      face_shape(ft, rs, N, dr, ds);
      double g1[3] = {0, 0, 0}, g2[3] = {0, 0, 0};
      for (int a = 0; a < nen; ++a)
        for (int i = 0; i < 3; ++i)
        {
          g1[i] += dr[a] * P[a][i];
          g2[i] += ds[a] * P[a][i];
        }
      const double fac = w * (gp_factor ? gp_factor[((size_t)face * ngp + g) * 3] : 1.0) * pressure;
      for (int a = 0; a < nen; ++a)
      {
        if (nodes[a] >= n_row_nodes) continue;
        const int64_t rb = node_rowbase[nodes[a]];
        const int rl = node_rowlen[nodes[a]];
        for (int b = 0; b < nen; ++b)
        {
          if (pos[a][b] < 0) continue;
          // D[dim][k] = d normal[dim] / d x_(b,k)
          const double D[3][3] = {{0.0, dr[b] * g2[2] - ds[b] * g1[2], ds[b] * g1[1] - dr[b] * g2[1]},
              {ds[b] * g1[2] - dr[b] * g2[2], 0.0, dr[b] * g2[0] - ds[b] * g1[0]},
              {dr[b] * g2[1] - ds[b] * g1[1], ds[b] * g1[0] - dr[b] * g2[0], 0.0}};
          for (int dim = 0; dim < 3; ++dim)
          {
            if (dbc && dbc[3 * nodes[a] + dim]) continue;
            for (int k = 0; k < 3; ++k)
              if (k != dim) red_add(K + rb + (int64_t)dim * rl + 3 * pos[a][b] + k, -N[a] * D[dim][k] * fac);
          }
        }
      }
    }
  }

  // Faces of nurbs27 elements (nurbs9, quad_9point): one thread per face, named by parent element and surface number like the
  // reference's surface element; its knot vectors are two of the parent's knot spans
  // (Knotvector::get_boundary_ele_and_parent_knots, 4C_fem_nurbs_discretization_knotvector.cpp:265-413).
  static __global__ void k_neumann_surface_nurbs(const double* __restrict__ X, const int* __restrict__ conn, const double* __restrict__ ele_knots,
      const double* __restrict__ node_weights, const int* __restrict__ parent, const int* __restrict__ surface, int n_faces, NeumannVal V,
      const double* __restrict__ gp_factor, int n_row_nodes, double* __restrict__ f, double* __restrict__ gp_coords)
  {
    const int face = blockIdx.x * blockDim.x + threadIdx.x;
    if (face >= n_faces) return;
    // eleNodeNumbering_nurbs27_surfaces, 4C_fem_general_utils_local_connectivity_matrices.hpp:250-253
    const unsigned char fn[6][9] = {{0, 1, 2, 3, 4, 5, 6, 7, 8}, {18, 19, 20, 21, 22, 23, 24, 25, 26}, {0, 1, 2, 9, 10, 11, 18, 19, 20},
        {6, 7, 8, 15, 16, 17, 24, 25, 26}, {2, 5, 8, 11, 14, 17, 20, 23, 26}, {0, 3, 6, 9, 12, 15, 18, 21, 24}};
    const int e = parent[face], sid = surface[face];
    const int d0 = sid < 4 ? 0 : 1, d1 = sid < 2 ? 1 : 2;
    double kn[12], w[9], P[9][3], fa[9][3];
    int nodes[9];
    for (int k = 0; k < 6; ++k)
    {
      kn[k] = ele_knots[(size_t)e * 18 + 6 * d0 + k];
      kn[6 + k] = ele_knots[(size_t)e * 18 + 6 * d1 + k];
    }
    for (int a = 0; a < 9; ++a)
    {
      nodes[a] = conn[(size_t)e * 27 + fn[sid][a]];
      w[a] = node_weights[nodes[a]];
      for (int i = 0; i < 3; ++i)
      {
        P[a][i] = X[3 * (size_t)nodes[a] + i];
        fa[a][i] = 0.0;
      }
    }
    for (int g = 0; g < 9; ++g)
    {
      double rs[2], gw, N[9], dN[18];
      face_gauss_point(QUAD9, g, rs, &gw);
      nurbs_funct_deriv<2>(rs, kn, w, N, dN);
      if (gp_coords)
      {
        for (int i = 0; i < 3; ++i)
        {
          double x = 0.0;
          for (int a = 0; a < 9; ++a) x += N[a] * P[a][i];
          gp_coords[((size_t)face * 9 + g) * 3 + i] = x;
        }
        continue;
      }
      double g1[3] = {0, 0, 0}, g2[3] = {0, 0, 0};
      for (int a = 0; a < 9; ++a)
        for (int i = 0; i < 3; ++i)
        {
          g1[i] += dN[2 * a] * P[a][i];
          g2[i] += dN[2 * a + 1] * P[a][i];
        }
      const double m11 = g1[0] * g1[0] + g1[1] * g1[1] + g1[2] * g1[2];
      const double m22 = g2[0] * g2[0] + g2[1] * g2[1] + g2[2] * g2[2];
      const double m12 = g1[0] * g2[0] + g1[1] * g2[1] + g1[2] * g2[2];
      const double detA = sqrt(m11 * m22 - m12 * m12);
      for (int dof = 0; dof < 3; ++dof)
      {
        if (!V.onoff[dof]) continue;
        const double functfac = gp_factor ? gp_factor[((size_t)face * 9 + g) * 3 + dof] : 1.0;
        const double fac = gw * detA * V.val[dof] * functfac;
        for (int a = 0; a < 9; ++a) fa[a][dof] += N[a] * fac;
      }
    }
    if (gp_coords) return;
    for (int a = 0; a < 9; ++a)
      if (nodes[a] < n_row_nodes)
        for (int dof = 0; dof < 3; ++dof)
          if (V.onoff[dof]) red_add(f + 3 * (size_t)nodes[a] + dof, fa[a][dof]);
  }

  // body loads on nurbs27 elements: N and dN of the Gauss point from the element's table (fc_nurbs.cuh)
  static __global__ void k_neumann_volume_nurbs(const double* __restrict__ X, const int* __restrict__ conn, const double* __restrict__ tab,
      const double* __restrict__ gpw, const int* __restrict__ ele_list, int n_list, NeumannVal V, const double* __restrict__ gp_factor,
      int n_row_nodes, double* __restrict__ f, double* __restrict__ gp_coords)
  {
    constexpr int NEN = 27, NGP = 27;
    const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (tid >= (long long)n_list * NGP) return;
    const int item = (int)(tid / NGP), g = (int)(tid % NGP);
    const int e = ele_list[item];
    const double* T = tab + (size_t)e * NurbsTab::stride;
    const double* N = T + NurbsTab::oN + g * NEN;
    const double* dN = T + (size_t)g * NEN * 3;
    double J[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0}, xg[3] = {0, 0, 0};
    for (int a = 0; a < NEN; ++a)
    {
      const int n = conn[(size_t)e * NEN + a];
      for (int i = 0; i < 3; ++i)
      {
        const double x = X[3 * (size_t)n + i];
        xg[i] += N[a] * x;
        for (int d = 0; d < 3; ++d) J[3 * d + i] += dN[3 * a + d] * x;
      }
    }
    if (gp_coords)
    {
      for (int i = 0; i < 3; ++i) gp_coords[((size_t)item * NGP + g) * 3 + i] = xg[i];
      return;
    }
    const double intfac = gpw[g] * det3(J);
    for (int dof = 0; dof < 3; ++dof)
    {
      if (!V.onoff[dof]) continue;
      const double functfac = gp_factor ? gp_factor[((size_t)item * NGP + g) * 3 + dof] : 1.0;
      const double vif = V.val[dof] * functfac * intfac;
      for (int a = 0; a < NEN; ++a)
      {
        const int n = conn[(size_t)e * NEN + a];
        if (n < n_row_nodes) red_add(f + 3 * (size_t)n + dof, N[a] * vif);
      }
    }
  }

  // one thread per (listed element, Gauss point)
  template <int CT>
  __global__ void k_neumann_volume(const double* __restrict__ X, const int* __restrict__ conn, const int* __restrict__ ele_list, int n_list, NeumannVal V,
      const double* __restrict__ gp_factor, int n_row_nodes, double* __restrict__ f, double* __restrict__ gp_coords)
  {
    constexpr int NEN = Cell<CT>::nen, NGP = Cell<CT>::ngp;
    const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (tid >= (long long)n_list * NGP) return;
    const int item = (int)(tid / NGP), g = (int)(tid % NGP);
    const int e = ele_list[item];
    double xi[3], w, N[NEN], dN[3 * NEN];
    gauss_point<CT>(g, xi, &w);
    shape_function<CT>(xi, N);
    shape_deriv1<CT>(xi, dN);
    double J[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0}, xg[3] = {0, 0, 0};
    for (int a = 0; a < NEN; ++a)
    {
      const int n = conn[(size_t)e * NEN + a];
      for (int i = 0; i < 3; ++i)
      {
        const double x = X[3 * (size_t)n + i];
        xg[i] += N[a] * x;
        for (int d = 0; d < 3; ++d) J[3 * d + i] += dN[3 * a + d] * x;
      }
    }
    if (gp_coords)
    {
      for (int i = 0; i < 3; ++i) gp_coords[((size_t)item * NGP + g) * 3 + i] = xg[i];
      return;
    }
    const double intfac = w * det3(J);
    for (int dof = 0; dof < 3; ++dof)
    {
      if (!V.onoff[dof]) continue;
      const double functfac = gp_factor ? gp_factor[((size_t)item * NGP + g) * 3 + dof] : 1.0;
      const double vif = V.val[dof] * functfac * intfac;
      for (int a = 0; a < NEN; ++a)
      {
        const int n = conn[(size_t)e * NEN + a];
        if (n < n_row_nodes) red_add(f + 3 * (size_t)n + dof, N[a] * vif);
      }
    }
  }
}  // namespace fc
