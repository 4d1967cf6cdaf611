// CPU ORACLE - test infrastructure only (see oracle/README.md).  Small dense helpers the reference takes from LAPACK (symmetric
// 3x3 eigenproblem, dense inverse) and the strain output types of struct_calc_stress:
//   assemble_strain_type_to_matrix_row     src/solid_3D_ele/4C_solid_3D_ele_calc_lib_io.hpp:113-153
//   green_lagrange_to_euler_almansi / green_lagrange_to_log_strain   src/solid_3D_ele/4C_solid_3D_ele_utils.cpp:35-84
// The oracle uses cyclic Jacobi rotations and Gauss-Jordan with partial pivoting - same results to rounding.
#pragma once
#include <cmath>
#include <utility>
#include <vector>

#include "fc_oracle_dense.hpp"

namespace fc_oracle
{
  // eigen-decomposition of a symmetric 3x3 (col-major): A = V diag(w) V^T
  inline void sym_eig3(const double* A, double* w, double* V)
  {
    double a[9];
    for (int i = 0; i < 9; ++i)
    {
      a[i] = A[i];
      V[i] = (i % 4 == 0) ? 1.0 : 0.0;
    }
    for (int sweep = 0; sweep < 60; ++sweep)
    {
      const double off = a[3] * a[3] + a[6] * a[6] + a[7] * a[7];
      const double diag = a[0] * a[0] + a[4] * a[4] + a[8] * a[8];
      if (off <= 1.0e-34 * diag || off == 0.0) break;
      for (int p = 0; p < 2; ++p)
        for (int q = p + 1; q < 3; ++q)
        {
          const double apq = a[p + 3 * q];
          if (apq == 0.0) continue;
          const double theta = (a[q + 3 * q] - a[p + 3 * p]) / (2.0 * apq);
          const double t = (theta >= 0.0 ? 1.0 : -1.0) / (std::fabs(theta) + std::sqrt(theta * theta + 1.0));
          const double c = 1.0 / std::sqrt(t * t + 1.0), s = t * c;
          for (int k = 0; k < 3; ++k)  // A <- A G
          {
            const double akp = a[k + 3 * p], akq = a[k + 3 * q];
            a[k + 3 * p] = c * akp - s * akq;
            a[k + 3 * q] = s * akp + c * akq;
          }
          for (int k = 0; k < 3; ++k)  // A <- G^T A
          {
            const double apk = a[p + 3 * k], aqk = a[q + 3 * k];
            a[p + 3 * k] = c * apk - s * aqk;
            a[q + 3 * k] = s * apk + c * aqk;
          }
          for (int k = 0; k < 3; ++k)
          {
            const double vkp = V[k + 3 * p], vkq = V[k + 3 * q];
            V[k + 3 * p] = c * vkp - s * vkq;
            V[k + 3 * q] = s * vkp + c * vkq;
          }
        }
    }
    for (int i = 0; i < 3; ++i) w[i] = a[i + 3 * i];
  }

  // f(A) = V diag(f(w)) V^T of a symmetric 3x3
  template <class Fn>
  inline void sym_function3(const double* A, Fn f, double* out)
  {
    double w[3], V[9];
    sym_eig3(A, w, V);
    for (int i = 0; i < 3; ++i) w[i] = f(w[i]);
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j) out[i + 3 * j] = V[i] * w[0] * V[j] + V[i + 3] * w[1] * V[j + 3] + V[i + 6] * w[2] * V[j + 6];
  }

  // dense inverse (col-major n x n) by Gauss-Jordan with partial pivoting; false if a pivot is zero
  inline bool invert_dense(int n, double* A)
  {
    std::vector<double> W((size_t)n * 2 * n, 0.0);  // row-major [A | I]
    for (int r = 0; r < n; ++r)
    {
      for (int c = 0; c < n; ++c) W[(size_t)r * 2 * n + c] = A[r + (size_t)n * c];
      W[(size_t)r * 2 * n + n + r] = 1.0;
    }
    for (int k = 0; k < n; ++k)
    {
      int piv = k;
      for (int r = k + 1; r < n; ++r)
        if (std::fabs(W[(size_t)r * 2 * n + k]) > std::fabs(W[(size_t)piv * 2 * n + k])) piv = r;
      if (W[(size_t)piv * 2 * n + k] == 0.0) return false;
      if (piv != k)
        for (int c = 0; c < 2 * n; ++c) std::swap(W[(size_t)k * 2 * n + c], W[(size_t)piv * 2 * n + c]);
      const double ip = 1.0 / W[(size_t)k * 2 * n + k];
      for (int c = 0; c < 2 * n; ++c) W[(size_t)k * 2 * n + c] *= ip;
      for (int r = 0; r < n; ++r)
      {
        if (r == k) continue;
        const double m = W[(size_t)r * 2 * n + k];
        if (m == 0.0) continue;
        for (int c = 0; c < 2 * n; ++c) W[(size_t)r * 2 * n + c] -= m * W[(size_t)k * 2 * n + c];
      }
    }
    for (int r = 0; r < n; ++r)
      for (int c = 0; c < n; ++c) A[r + (size_t)n * c] = W[(size_t)r * 2 * n + n + c];
    return true;
  }

  // strain output of one Gauss point from the Green-Lagrange strain (strain-like Voigt) and the deformation gradient
  // (col-major); out in stress-like Voigt (xx yy zz xy yz xz)
  enum StrainType
  {
    STRAIN_GL = 0,
    STRAIN_EA = 1,
    STRAIN_LOG = 2
  };
  inline void strain_output(int type, const double* E, const double* F, double* out)
  {
    const double Em[9] = {E[0], 0.5 * E[3], 0.5 * E[5], 0.5 * E[3], E[1], 0.5 * E[4], 0.5 * E[5], 0.5 * E[4], E[2]};
    double R[9];
    if (type == STRAIN_GL)
      for (int i = 0; i < 9; ++i) R[i] = Em[i];
    else if (type == STRAIN_EA)
    {
      double iF[9], t[9];
      invert3x3(iF, F);
      multiply_tn<3, 3, 3>(t, iF, Em);  // F^-T E
      multiply_nn<3, 3, 3>(R, t, iF);   // F^-T E F^-1
    }
    else
      sym_function3(Em, [](double e) { return std::log(std::sqrt(2.0 * e + 1.0)); }, R);
    out[0] = R[0];
    out[1] = R[4];
    out[2] = R[8];
    out[3] = 0.5 * (R[3] + R[1]);
    out[4] = 0.5 * (R[7] + R[5]);
    out[5] = 0.5 * (R[6] + R[2]);
  }
}  // namespace fc_oracle
