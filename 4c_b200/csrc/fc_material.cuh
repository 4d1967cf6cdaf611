// Device material laws of the path: PK2 stress (stress-like Voigt xx,yy,zz,xy,yz,zx) and material
// tangent (6x6, symmetric, Voigt stress/stress notation as consumed by B^T C B with a strain-like B).
//   St.Venant-Kirchhoff   src/mat/4C_mat_stvenantkirchhoff.cpp:115-145,169-177
//   ElastHyper toolbox    src/mat/4C_mat_elasthyper_service.cpp:19-78 (driver), :162-215 (isotropic
//                         stress + cmat from gamma/delta), :413-431 (gamma/delta)
//   CoupNeoHooke summand  src/mat/elast/4C_mat_elast_coupneohooke.cpp:17-25,60-77
//   summand list          the loops over potsum_ of 4C_mat_elasthyper_service.cpp:90-104 (isotropic principal
//                         derivatives) and :386-396 (anisotropic principal stress + cmat)
//   CoupSVK summand       src/mat/elast/4C_mat_elast_coupSaintVenantKirchhoff.cpp:15-29,47-57
//   CoupAnisoExpo summand src/mat/elast/4C_mat_elast_coupanisoexpobase.cpp:157-185, one fibre per element (FIBER1,
//                         normalised: src/mat/4C_mat_anisotropy_utils.cpp:19-44), structural tensor a a^T
// Written in tensor form (no Voigt scale-factor tables): the same algebra as the reference, a
// different evaluation order.
#pragma once
#include "fc_fem.cuh"

namespace fc
{
  enum
  {
    MAT_SVK = 0,
    MAT_NEOHOOKE = 1,   // MAT_ElastHyper with one CoupNeoHooke summand (young, nue): the closed-form kernels
    MAT_ELASTHYPER = 2  // MAT_ElastHyper with the summand list below: generic kernel
  };
  enum
  {
    SUM_COUPNEOHOOKE = 0,  // par = YOUNG, NUE
    SUM_COUPSVK = 1,       // par = YOUNG, NUE
    SUM_COUPANISOEXPO = 2  // par = K1, K2, K1COMP, K2COMP
  };
  constexpr int max_summands = 4;

  struct MatParams
  {
    int kind;
    double young, nue;
    int n_summands;
    int sum_kind[max_summands];
    double sum_par[max_summands][4];
  };

  // symmetric 3x3 stored as 6: [00,11,22,01,12,02]
  FC_HD double sym_det(const double* c)
  {
    return c[0] * (c[1] * c[2] - c[4] * c[4]) - c[3] * (c[3] * c[2] - c[4] * c[5]) +
           c[5] * (c[3] * c[4] - c[1] * c[5]);
  }
  FC_HD void sym_inv(const double* c, double det, double* ic)
  {
    const double id = 1.0 / det;
    ic[0] = (c[1] * c[2] - c[4] * c[4]) * id;
    ic[1] = (c[0] * c[2] - c[5] * c[5]) * id;
    ic[2] = (c[0] * c[1] - c[3] * c[3]) * id;
    ic[3] = (c[5] * c[4] - c[3] * c[2]) * id;
    ic[4] = (c[3] * c[5] - c[0] * c[4]) * id;
    ic[5] = (c[3] * c[4] - c[5] * c[1]) * id;
  }

  // Invariant derivatives of ELAST_CoupNeoHooke: Psi = c (I1-3) + c/beta (I3^-beta - 1)
  FC_HD void coupneohooke_derivs(double young, double nue, const double* prinv, double* dPI, double* ddPII)
  {
    const double c = young / (4.0 * (1.0 + nue));
    const double beta = nue / (1.0 - 2.0 * nue);
    dPI[0] += c;
    if (prinv[2] > 0)
    {
      const double p = exp(log(prinv[2]) * (-beta - 1.0));
      dPI[2] -= c * p;
      ddPII[2] += c * (beta + 1.0) * p / prinv[2];
    }
    else
    {
      dPI[2] = ddPII[2] = nan("");
    }
  }

  // Invariant derivatives of ELAST_CoupSVK: Psi = (mue/4 + lambda/8) I1^2 - (3 lambda/4 + mue/2) I1 - mue/2 I2 + const
  FC_HD void coupsvk_derivs(double young, double nue, const double* prinv, double* dPI, double* ddPII)
  {
    const double lambda = nue == 0.5 ? 0.0 : young * nue / ((1.0 + nue) * (1.0 - 2.0 * nue));
    const double mue = young / (2.0 * (1.0 + nue));
    dPI[0] += (0.5 * mue + 0.25 * lambda) * prinv[0] - 0.75 * lambda - 0.5 * mue;
    dPI[1] -= 0.5 * mue;
    ddPII[0] += 0.5 * mue + 0.25 * lambda;
  }

  FC_HD void gamma_delta(const double* prinv, const double* dPI, const double* ddPII, double* gamma, double* delta)
  {
    gamma[0] = 2.0 * (dPI[0] + prinv[0] * dPI[1]);
    gamma[1] = -2.0 * dPI[1];
    gamma[2] = 2.0 * prinv[2] * dPI[2];
    delta[0] = 4.0 * (ddPII[0] + 2.0 * prinv[0] * ddPII[5] + dPI[1] + prinv[0] * prinv[0] * ddPII[1]);
    delta[1] = -4.0 * (ddPII[5] + prinv[0] * ddPII[1]);
    delta[2] = 4.0 * (prinv[2] * ddPII[4] + prinv[0] * prinv[2] * ddPII[3]);
    delta[3] = 4.0 * ddPII[1];
    delta[4] = -4.0 * prinv[2] * ddPII[3];
    delta[5] = 4.0 * (prinv[2] * dPI[2] + prinv[2] * prinv[2] * ddPII[2]);
    delta[6] = -4.0 * prinv[2] * dPI[2];
    delta[7] = -4.0 * dPI[1];
  }

  // Voigt index -> tensor index pair
  FC_HD int vi(int v) { return v < 3 ? v : (v == 3 ? 0 : (v == 4 ? 1 : 0)); }
  FC_HD int vj(int v) { return v < 3 ? v : (v == 3 ? 1 : 2); }
  // symmetric tensor index pair -> storage index in [00,11,22,01,12,02]
  FC_HD int sidx(int i, int j)
  {
    if (i == j) return i;
    const int s = i + j;  // 1 -> 01, 3 -> 12, 2 -> 02
    return s == 1 ? 3 : (s == 3 ? 4 : 5);
  }

  // C: right Cauchy-Green as symmetric-6 (true tensor components, no doubled shear).
  // S[6], cmat[36] (cmat[6*r+c], symmetric).
  // fiber [3]: unit fibre direction of the element (anisotropic summands only)
  FC_HD void material_evaluate(const MatParams& mp, const double* C, double* S, double* cmat, const double* fiber = nullptr)
  {
    if (mp.kind == MAT_SVK)
    {
      const double mfac = mp.young / ((1.0 + mp.nue) * (1.0 - 2.0 * mp.nue));
      const double c00 = mfac * (1.0 - mp.nue), c01 = mfac * mp.nue, c33 = mfac * 0.5 * (1.0 - 2.0 * mp.nue);
      for (int i = 0; i < 36; ++i) cmat[i] = 0.0;
      for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) cmat[6 * i + j] = (i == j) ? c00 : c01;
      for (int i = 3; i < 6; ++i) cmat[6 * i + i] = c33;
      // strain-like Green-Lagrange: E_ii = (C_ii-1)/2, shear entries doubled -> 2*E_ij = C_ij
      const double E[6] = {0.5 * (C[0] - 1.0), 0.5 * (C[1] - 1.0), 0.5 * (C[2] - 1.0), C[3], C[4], C[5]};
      for (int i = 0; i < 3; ++i) S[i] = c00 * E[i] + c01 * (E[(i + 1) % 3] + E[(i + 2) % 3]);
      for (int i = 3; i < 6; ++i) S[i] = c33 * E[i];
      return;
    }
    // MAT_ElastHyper, isotropic principal-invariant summands
    double prinv[3], iC[6];
    prinv[0] = C[0] + C[1] + C[2];
    prinv[1] = 0.5 * (prinv[0] * prinv[0] - C[0] * C[0] - C[1] * C[1] - C[2] * C[2]) - C[3] * C[3] -
               C[4] * C[4] - C[5] * C[5];
    prinv[2] = sym_det(C);
    sym_inv(C, prinv[2], iC);
    double dPI[3] = {0, 0, 0}, ddPII[6] = {0, 0, 0, 0, 0, 0};
    if (mp.kind == MAT_NEOHOOKE)
      coupneohooke_derivs(mp.young, mp.nue, prinv, dPI, ddPII);
    else
      for (int k = 0; k < mp.n_summands; ++k)
      {
        if (mp.sum_kind[k] == SUM_COUPNEOHOOKE) coupneohooke_derivs(mp.sum_par[k][0], mp.sum_par[k][1], prinv, dPI, ddPII);
        if (mp.sum_kind[k] == SUM_COUPSVK) coupsvk_derivs(mp.sum_par[k][0], mp.sum_par[k][1], prinv, dPI, ddPII);
      }
    double gamma[3], delta[8];
    gamma_delta(prinv, dPI, ddPII, gamma, delta);
    const double id2[6] = {1, 1, 1, 0, 0, 0};
    for (int v = 0; v < 6; ++v) S[v] = gamma[0] * id2[v] + gamma[1] * C[v] + gamma[2] * iC[v];
    for (int r = 0; r < 6; ++r)
      for (int c = r; c < 6; ++c)
      {
        const int i = vi(r), j = vj(r), k = vi(c), l = vj(c);
        double v = delta[0] * id2[r] * id2[c] + delta[1] * (id2[r] * C[c] + C[r] * id2[c]) +
                   delta[2] * (id2[r] * iC[c] + iC[r] * id2[c]) + delta[3] * C[r] * C[c] +
                   delta[4] * (C[r] * iC[c] + iC[r] * C[c]) + delta[5] * iC[r] * iC[c];
        v += delta[6] * 0.5 * (iC[sidx(i, k)] * iC[sidx(j, l)] + iC[sidx(i, l)] * iC[sidx(j, k)]);
        if (r == c) v += delta[7] * (r < 3 ? 1.0 : 0.5);
        cmat[6 * r + c] = v;
        cmat[6 * c + r] = v;
      }
    if (mp.kind != MAT_ELASTHYPER || !fiber) return;
    // anisotropic principal summands: S += 2 dPsi/dI4 A, cmat += 4 d2Psi/dI4^2 A x A, I4 = a . C a
    for (int k = 0; k < mp.n_summands; ++k)
      if (mp.sum_kind[k] == SUM_COUPANISOEXPO)
      {
        const double A[6] = {fiber[0] * fiber[0], fiber[1] * fiber[1], fiber[2] * fiber[2], fiber[0] * fiber[1], fiber[1] * fiber[2], fiber[0] * fiber[2]};
        const double I4 = A[0] * C[0] + A[1] * C[1] + A[2] * C[2] + 2.0 * (A[3] * C[3] + A[4] * C[4] + A[5] * C[5]);
        const double k1 = I4 < 1.0 ? mp.sum_par[k][2] : mp.sum_par[k][0], k2 = I4 < 1.0 ? mp.sum_par[k][3] : mp.sum_par[k][1];
        const double d = I4 - 1.0, ex = exp(k2 * d * d);
        const double gam = 2.0 * k1 * d * ex, del = 4.0 * (1.0 + 2.0 * k2 * d * d) * k1 * ex;
        for (int r = 0; r < 6; ++r)
        {
          S[r] += gam * A[r];
          for (int c = 0; c < 6; ++c) cmat[6 * r + c] += del * A[r] * A[c];
        }
      }
  }
}  // namespace fc
