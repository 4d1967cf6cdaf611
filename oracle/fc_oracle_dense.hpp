// ORACLE — TEST INFRASTRUCTURE ONLY. Never linked into or called by the product library.
//
// CPU restatement of the fixed-size dense linear algebra and Voigt helpers of 4C that the solid
// element hot path uses.  Every function cites the reference file:line it follows (paths relative
// to /root/reference).  Summation order and operation order follow the reference so that results
// agree to the last bits when compiled with -ffp-contract=off.
//
// Parity status: the multiply / invert / determinant / Voigt determinant+invariants functions are
// pinned bit-exactly against the reference's own header compiled in oracle/_ref (see
// oracle/ref_dense_harness.cpp and tests/test_oracle_ref.py).  inverse_tensor / holzapfel product
// are pinned by the reference's unit-test golden values (tests/test_oracle_golden.py).
#pragma once
#include <cmath>
#include <cstring>

namespace fc_oracle
{
  // Column-major fixed-size matrix, same storage convention as Core::LinAlg::Matrix
  // (src/core/linalg/src/dense/4C_linalg_fixedsizematrix.hpp: data_[r + c*rows]).
  template <int R, int C>
  struct M
  {
    double v[R * C];
    double& operator()(int r, int c = 0) { return v[r + c * R]; }
    const double& operator()(int r, int c = 0) const { return v[r + c * R]; }
    void clear() { std::memset(v, 0, sizeof(v)); }
  };

  // out(IxK) = left(IxJ) * right(JxK)      4C_linalg_fixedsizematrix.hpp:905-928
  // first product initialises, remaining ones are added in ascending inner index.
  template <int I, int J, int K>
  inline void multiply_nn(double* out, const double* left, const double* right)
  {
    for (int c1 = 0; c1 < K; ++c1)
      for (int c2 = 0; c2 < I; ++c2)
      {
        double tmp = left[c2] * right[c1 * J];
        for (int c3 = 1; c3 < J; ++c3) tmp += left[c2 + c3 * I] * right[c1 * J + c3];
        out[c2 + c1 * I] = tmp;
      }
  }

  // out(IxK) = left(IxJ) * right(KxJ)^T    4C_linalg_fixedsizematrix.hpp:930-955
  template <int I, int J, int K>
  inline void multiply_nt(double* out, const double* left, const double* right)
  {
    for (int c1 = 0; c1 < K; ++c1)
      for (int c2 = 0; c2 < I; ++c2)
      {
        double tmp = left[c2] * right[c1];
        for (int c3 = 1; c3 < J; ++c3) tmp += left[c2 + c3 * I] * right[c1 + c3 * K];
        out[c2 + c1 * I] = tmp;
      }
  }

  // out(IxK) = left(JxI)^T * right(JxK)    4C_linalg_fixedsizematrix.hpp:957-985
  template <int I, int J, int K>
  inline void multiply_tn(double* out, const double* left, const double* right)
  {
    for (int c1 = 0; c1 < K; ++c1)
      for (int c2 = 0; c2 < I; ++c2)
      {
        double tmp = left[c2 * J] * right[c1 * J];
        for (int c3 = 1; c3 < J; ++c3) tmp += left[c2 * J + c3] * right[c1 * J + c3];
        out[c2 + c1 * I] = tmp;
      }
  }

  // out = infac * left * right^T + outfac * out   (scaled "update" overloads of the above:
  // 4C_linalg_fixedsizematrix.hpp, multiply_nt(outfac-variants): out = out*outfac + infac*tmp)
  template <int I, int J, int K>
  inline void multiply_nt_scaled(
      double infac, double* out, const double* left, const double* right, double outfac)
  {
    for (int c1 = 0; c1 < K; ++c1)
      for (int c2 = 0; c2 < I; ++c2)
      {
        double tmp = left[c2] * right[c1];
        for (int c3 = 1; c3 < J; ++c3) tmp += left[c2 + c3 * I] * right[c1 + c3 * K];
        double& o = out[c2 + c1 * I];
        o = (outfac == 0.0) ? infac * tmp : o * outfac + infac * tmp;
      }
  }

  template <int I, int J, int K>
  inline void multiply_tn_scaled(
      double infac, double* out, const double* left, const double* right, double outfac)
  {
    for (int c1 = 0; c1 < K; ++c1)
      for (int c2 = 0; c2 < I; ++c2)
      {
        double tmp = left[c2 * J] * right[c1 * J];
        for (int c3 = 1; c3 < J; ++c3) tmp += left[c2 * J + c3] * right[c1 * J + c3];
        double& o = out[c2 + c1 * I];
        o = (outfac == 0.0) ? infac * tmp : o * outfac + infac * tmp;
      }
  }

  // closed-form 3x3 inverse, returns determinant.   4C_linalg_fixedsizematrix.hpp:1304-1331
  // returns NaN determinant marker via *singular when det == 0 (reference throws).
  inline double invert3x3(double* out, const double* in, bool* singular = nullptr)
  {
    out[0] = in[4] * in[8] - in[5] * in[7];
    out[1] = in[2] * in[7] - in[1] * in[8];
    out[2] = in[1] * in[5] - in[2] * in[4];
    const double det = in[0] * out[0] + in[3] * out[1] + in[6] * out[2];
    if (det == 0.0 && singular) *singular = true;
    const double invdet = 1.0 / det;
    out[0] *= invdet;
    out[1] *= invdet;
    out[2] *= invdet;
    out[3] = invdet * (in[5] * in[6] - in[3] * in[8]);
    out[4] = invdet * (in[0] * in[8] - in[2] * in[6]);
    out[5] = invdet * (in[2] * in[3] - in[0] * in[5]);
    out[6] = invdet * (in[3] * in[7] - in[4] * in[6]);
    out[7] = invdet * (in[1] * in[6] - in[0] * in[7]);
    out[8] = invdet * (in[0] * in[4] - in[1] * in[3]);
    return det;
  }

  // 3x3 determinant.   4C_linalg_fixedsizematrix.hpp:1441-1455
  inline double determinant3x3(const double* m)
  {
    return m[0] * (m[4] * m[8] - m[5] * m[7]) + m[3] * (m[2] * m[7] - m[1] * m[8]) +
           m[6] * (m[1] * m[5] - m[2] * m[4]);
  }

  // ---- Voigt helpers (order xx,yy,zz,xy,yz,zx) ------------------------------------------------
  // strain-like notation carries doubled shear:
  // 4C_linalg_fixedsizematrix_voigt_notation.hpp:193-203 (unscale 0.5 / scale 2 on entries 3..5)
  inline double unscale_strain(int i) { return i < 3 ? 1.0 : 0.5; }
  inline double scale_strain(int i) { return i < 3 ? 1.0 : 2.0; }

  // triple_entry_product for strain-like tensors   ...voigt_notation.hpp:214-219
  inline double tep_strain(const double* t, int i, int j, int k)
  {
    return t[i] * unscale_strain(i) * t[j] * unscale_strain(j) * t[k] * unscale_strain(k);
  }
  inline double tep_stress(const double* t, int i, int j, int k) { return t[i] * 1.0 * t[j] * 1.0 * t[k] * 1.0; }

  // determinant of a Voigt tensor   ...voigt_notation.hpp:79-84
  inline double voigt_determinant_strain(const double* t)
  {
    return tep_strain(t, 0, 1, 2) + 2 * tep_strain(t, 3, 4, 5) - tep_strain(t, 1, 5, 5) -
           tep_strain(t, 2, 3, 3) - tep_strain(t, 0, 4, 4);
  }
  inline double voigt_determinant_stress(const double* t)
  {
    return tep_stress(t, 0, 1, 2) + 2 * tep_stress(t, 3, 4, 5) - tep_stress(t, 1, 5, 5) -
           tep_stress(t, 2, 3, 3) - tep_stress(t, 0, 4, 4);
  }

  // principal invariants of a strain-like Voigt tensor   ...voigt_notation.hpp:90-104
  inline void voigt_invariants_principal_strain(double* prinv, const double* t)
  {
    prinv[0] = t[0] + t[1] + t[2];
    prinv[1] = 0.5 * (prinv[0] * prinv[0] - t[0] * t[0] - t[1] * t[1] - t[2] * t[2]) -
               t[3] * t[3] * unscale_strain(3) * unscale_strain(3) -
               t[4] * t[4] * unscale_strain(4) * unscale_strain(4) -
               t[5] * t[5] * unscale_strain(5) * unscale_strain(5);
    prinv[2] = voigt_determinant_strain(t);
  }
  inline void voigt_invariants_principal_stress(double* prinv, const double* t)
  {
    prinv[0] = t[0] + t[1] + t[2];
    prinv[1] = 0.5 * (prinv[0] * prinv[0] - t[0] * t[0] - t[1] * t[1] - t[2] * t[2]) -
               t[3] * t[3] * 1.0 * 1.0 - t[4] * t[4] * 1.0 * 1.0 - t[5] * t[5] * 1.0 * 1.0;
    prinv[2] = voigt_determinant_stress(t);
  }

  // inverse of a symmetric tensor in Voigt notation
  // 4C_linalg_fixedsizematrix_voigt_notation.cpp:188-207 (strain-like: us = 0.5, sc = 2)
  template <bool strain>
  inline void voigt_inverse_tensor(const double* t, double* ti)
  {
    auto us = [](int i) { return strain ? unscale_strain(i) : 1.0; };
    auto sc = [](int i) { return strain ? scale_strain(i) : 1.0; };
    const double det = strain ? voigt_determinant_strain(t) : voigt_determinant_stress(t);
    ti[0] = (t[1] * t[2] - us(4) * us(4) * t[4] * t[4]) / det * sc(0);
    ti[1] = (t[0] * t[2] - us(5) * us(5) * t[5] * t[5]) / det * sc(1);
    ti[2] = (t[0] * t[1] - us(3) * us(3) * t[3] * t[3]) / det * sc(2);
    ti[3] = (us(5) * us(4) * t[5] * t[4] - us(3) * us(2) * t[3] * t[2]) / det * sc(3);
    ti[4] = (us(3) * us(5) * t[3] * t[5] - us(0) * us(4) * t[0] * t[4]) / det * sc(4);
    ti[5] = (us(3) * us(4) * t[3] * t[4] - us(5) * us(1) * t[5] * t[1]) / det * sc(5);
  }

  // strain-like -> stress-like   ...voigt_notation.cpp:212-216
  inline void voigt_strain_to_stress_like(const double* in, double* out)
  {
    for (int i = 0; i < 6; ++i) out[i] = unscale_strain(i) * in[i];
  }

  // 3x3 (col-major) -> strain-like Voigt vector   ...voigt_notation.cpp:238-245
  inline void voigt_matrix_to_vector_strain(const double* m, double* vout)
  {
    for (int i = 0; i < 3; ++i) vout[i] = m[i + 3 * i];
    vout[3] = 0.5 * scale_strain(3) * (m[0 + 3 * 1] + m[1 + 3 * 0]);
    vout[4] = 0.5 * scale_strain(4) * (m[1 + 3 * 2] + m[2 + 3 * 1]);
    vout[5] = 0.5 * scale_strain(5) * (m[0 + 3 * 2] + m[2 + 3 * 0]);
  }

  // cmat += scalar * (invc (.) invc)  "Holzapfel product", invc stress-like, cmat 6x6 col-major
  // 4C_linalg_fixedsizematrix_tensor_products.cpp:264-314.  Written from the index rule
  //   (A(.)A)_{(ij)(kl)} = 1/2 (A_ik A_jl + A_il A_jk)
  // but evaluated term-for-term like the reference: normal/normal and normal/shear entries are a
  // single product (the two halves coincide), shear/shear entries are 0.5*(p1 + p2).
  inline void add_holzapfel_product(double* cmat, const double* invc, double scalar)
  {
    // Voigt index -> tensor index pair
    static const int vi[6] = {0, 1, 2, 0, 1, 0};
    static const int vj[6] = {0, 1, 2, 1, 2, 2};
    // tensor (i,j) -> Voigt index
    static const int tv[3][3] = {{0, 3, 5}, {3, 1, 4}, {5, 4, 2}};
    for (int r = 0; r < 6; ++r)
      for (int c = 0; c < 6; ++c)
      {
        const int i = vi[r], j = vj[r], k = vi[c], l = vj[c];
        double val;
        if (r < 3 || c < 3)
        {
          // halves coincide: A_ik A_jl with the reference's factor order (see table in the .cpp)
          const double a = invc[tv[i][k]], b = invc[tv[j][l]];
          val = scalar * a * b;
        }
        else
        {
          val = scalar * 0.5 *
                (invc[tv[i][k]] * invc[tv[j][l]] + invc[tv[i][l]] * invc[tv[j][k]]);
        }
        cmat[r + 6 * c] += val;
      }
  }
}  // namespace fc_oracle
