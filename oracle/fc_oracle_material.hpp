// ORACLE — TEST INFRASTRUCTURE ONLY. Never linked into or called by the product library.
//
// Restatement of Mat::So3Material::evaluate for the two laws on the path
// (paths relative to /root/reference):
//   St.Venant-Kirchhoff      src/mat/4C_mat_stvenantkirchhoff.cpp:115-145 (fill_cmat), :169-177
//   MAT_ElastHyper + ELAST_CoupNeoHooke
//        src/mat/4C_mat_elasthyper.cpp:337-346  -> src/mat/4C_mat_elasthyper_service.cpp:19-78
//        C = 2E + I                 ...service.cpp:80-88
//        isotropic stress + cmat    ...service.cpp:162-215
//        gamma / delta              ...service.cpp:413-431
//        summand derivatives        src/mat/elast/4C_mat_elast_coupneohooke.cpp:17-25, :60-77
//   MAT_ElastHyper with a LIST of summands: the loop over potsum_ of ...service.cpp:90-104 and :386-396
//        ELAST_CoupSVK        src/mat/elast/4C_mat_elast_coupSaintVenantKirchhoff.cpp:15-29 (lambda, mue), :47-57
//        ELAST_CoupAnisoExpo  src/mat/elast/4C_mat_elast_coupanisoexpobase.cpp:157-185 (add_stress_aniso_principal),
//                             fibre from the element line, normalised (src/mat/4C_mat_anisotropy_utils.cpp:19-44),
//                             structural tensor A = a a^T, scalar product 1 (4C_mat_elast_coupanisoexpo.cpp:25-28)
#pragma once
#include <cmath>
#include <limits>

#include "fc_oracle_dense.hpp"

namespace fc_oracle
{
  enum MaterialKind
  {
    MAT_SVK = 0,
    MAT_NEOHOOKE = 1,   // MAT_ElastHyper with one ELAST_CoupNeoHooke summand (young, nue)
    MAT_ELASTHYPER = 2  // MAT_ElastHyper with the summand list below
  };
  enum SummandKind
  {
    SUM_COUPNEOHOOKE = 0,  // par = YOUNG, NUE
    SUM_COUPSVK = 1,       // par = YOUNG, NUE
    SUM_COUPANISOEXPO = 2  // par = K1, K2, K1COMP, K2COMP (fibre: FIBER1 of the element)
  };

  struct MaterialParams
  {
    int kind;
    double young;
    double nue;
    int n_summands = 0;
    int sum_kind[4] = {0, 0, 0, 0};
    double sum_par[4][4] = {{0, 0, 0, 0}, {0, 0, 0, 0}, {0, 0, 0, 0}, {0, 0, 0, 0}};
  };

  // Holzapfel gamma/delta coefficients from invariant derivatives.  ...service.cpp:413-431
  inline void calculate_gamma_delta(
      double* gamma, double* delta, const double* prinv, const double* dPI, const double* ddPII)
  {
    gamma[0] = 2. * (dPI[0] + prinv[0] * dPI[1]);
    gamma[1] = -2. * dPI[1];
    gamma[2] = 2. * prinv[2] * dPI[2];

    delta[0] = 4. * (ddPII[0] + 2. * prinv[0] * ddPII[5] + dPI[1] + prinv[0] * prinv[0] * ddPII[1]);
    delta[1] = -4. * (ddPII[5] + prinv[0] * ddPII[1]);
    delta[2] = 4. * (prinv[2] * ddPII[4] + prinv[0] * prinv[2] * ddPII[3]);
    delta[3] = 4. * ddPII[1];
    delta[4] = -4. * prinv[2] * ddPII[3];
    delta[5] = 4. * (prinv[2] * dPI[2] + prinv[2] * prinv[2] * ddPII[2]);
    delta[6] = -4. * prinv[2] * dPI[2];
    delta[7] = -4. * dPI[1];
  }

  // ELAST_CoupNeoHooke: Psi = c (I1-3) + c/beta (I3^-beta - 1).  coupneohooke.cpp:60-77
  inline void coupneohooke_add_derivatives_principal(
      double young, double nue, double* dPI, double* ddPII, const double* prinv)
  {
    const double c = young / (4.0 * (1.0 + nue));  // :22-24
    const double beta = nue / (1.0 - 2.0 * nue);
    dPI[0] += c;
    if (prinv[2] > 0)
    {
      const double prinv2_to_beta_m1 = std::exp(std::log(prinv[2]) * (-beta - 1.));
      dPI[2] -= c * prinv2_to_beta_m1;
      ddPII[2] += c * (beta + 1.) * prinv2_to_beta_m1 / prinv[2];
    }
    else
      dPI[2] = ddPII[2] = std::numeric_limits<double>::quiet_NaN();
  }

  // ELAST_CoupSVK.  coupSaintVenantKirchhoff.cpp:15-29 (parameters), :47-57 (derivatives)
  inline void coupsvk_add_derivatives_principal(double c1, double c2, double* dPI, double* ddPII, const double* prinv)
  {
    const double lambda = (c2 == 0.5) ? 0.0 : c1 * c2 / ((1.0 + c2) * (1.0 - 2.0 * c2));
    const double mue = c1 / (2.0 * (1.0 + c2));
    dPI[0] += (0.5 * mue + 0.25 * lambda) * prinv[0] - 0.75 * lambda - 0.5 * mue;
    dPI[1] -= 0.5 * mue;
    ddPII[0] += 0.5 * mue + 0.25 * lambda;
  }

  // rank-one update cmat += fac * a b^T   (multiply_nt(fac, a, b, 1.0) with 6x1 operands)
  inline void add_dyad6(double* cmat, double fac, const double* a, const double* b)
  {
    for (int c1 = 0; c1 < 6; ++c1)
      for (int c2 = 0; c2 < 6; ++c2)
      {
        const double tmp = a[c2] * b[c1];
        cmat[c2 + 6 * c1] = cmat[c2 + 6 * c1] * 1.0 + fac * tmp;
      }
  }

  // elast_hyper_add_isotropic_stress_cmat   ...service.cpp:162-215
  inline void elast_hyper_add_isotropic_stress_cmat(double* S, double* cmat, const double* C_strain,
      const double* iC_strain, const double* prinv, const double* dPI, const double* ddPII)
  {
    double gamma[3], delta[8];
    const double id2[6] = {1, 1, 1, 0, 0, 0};
    double C_stress[6], iC_stress[6];
    voigt_strain_to_stress_like(C_strain, C_stress);
    voigt_strain_to_stress_like(iC_strain, iC_stress);
    calculate_gamma_delta(gamma, delta, prinv, dPI, ddPII);

    for (int i = 0; i < 6; ++i) S[i] = S[i] * 1.0 + gamma[0] * id2[i];
    for (int i = 0; i < 6; ++i) S[i] = S[i] * 1.0 + gamma[1] * C_stress[i];
    for (int i = 0; i < 6; ++i) S[i] = S[i] * 1.0 + gamma[2] * iC_stress[i];

    add_dyad6(cmat, delta[0], id2, id2);
    add_dyad6(cmat, delta[1], id2, C_stress);
    add_dyad6(cmat, delta[1], C_stress, id2);
    add_dyad6(cmat, delta[2], id2, iC_stress);
    add_dyad6(cmat, delta[2], iC_stress, id2);
    add_dyad6(cmat, delta[3], C_stress, C_stress);
    add_dyad6(cmat, delta[4], C_stress, iC_stress);
    add_dyad6(cmat, delta[4], iC_stress, C_stress);
    add_dyad6(cmat, delta[5], iC_stress, iC_stress);
    add_holzapfel_product(cmat, iC_stress, delta[6]);
    // delta7 * Id4^# (stress/stress notation: diag 1,1,1,.5,.5,.5)  voigt_notation.cpp:42-52
    for (int i = 0; i < 6; ++i) cmat[i + 6 * i] = cmat[i + 6 * i] * 1.0 + delta[7] * (i < 3 ? 1.0 : 0.5);
  }

  // CoupAnisoExpoBase::add_stress_aniso_principal.  coupanisoexpobase.cpp:157-185; rcg = C strain-like,
  // fiber = FIBER1 of the element as read (normalised here like read_anisotropy_fiber)
  inline void coupanisoexpo_add_stress_aniso_principal(
      double k1_, double k2_, double k1comp_, double k2comp_, const double* fiber, const double* rcg, double* cmat, double* stress)
  {
    double f1norm = 0.;
    for (int i = 0; i < 3; ++i) f1norm += fiber[i] * fiber[i];
    f1norm = std::sqrt(f1norm);
    double a[3];
    for (int i = 0; i < 3; ++i) a[i] = fiber[i] / f1norm;
    // structural tensor in stress-like Voigt notation
    const double A_stress[6] = {a[0] * a[0], a[1] * a[1], a[2] * a[2], a[0] * a[1], a[1] * a[2], a[0] * a[2]};
    double I4 = 0.0;
    for (int i = 0; i < 6; ++i) I4 += A_stress[i] * rcg[i];
    const double scalarProduct = 1.0;
    double k1, k2;
    if (I4 < 1.0)
    {
      k1 = k1comp_;
      k2 = k2comp_;
    }
    else
    {
      k1 = k1_;
      k2 = k2_;
    }
    const double gamma = 2. * (k1 * (I4 - scalarProduct) * std::exp(k2 * std::pow((I4 - scalarProduct), 2)));
    for (int i = 0; i < 6; ++i) stress[i] = stress[i] * 1.0 + gamma * A_stress[i];
    const double delta = 2. * (1. + 2. * k2 * std::pow((I4 - scalarProduct), 2)) * 2. * k1 * std::exp(k2 * std::pow((I4 - scalarProduct), 2));
    add_dyad6(cmat, delta, A_stress, A_stress);
  }

  // So3Material::evaluate: glstrain (strain-like Voigt) -> stress (stress-like), cmat 6x6 col-major.  fiber [3] = the
  // element's FIBER1 (needed by anisotropic summands only)
  inline void material_evaluate(
      const MaterialParams& mp, const double* defgrd, const double* glstrain, double* S, double* cmat, const double* fiber = nullptr)
  {
    (void)defgrd;
    if (mp.kind == MAT_SVK)
    {
      const double Emod = mp.young, nu = mp.nue;
      const double mfac = Emod / ((1.0 + nu) * (1.0 - 2.0 * nu));
      for (int i = 0; i < 36; ++i) cmat[i] = 0.0;
      for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) cmat[i + 6 * j] = (i == j) ? mfac * (1.0 - nu) : mfac * nu;
      for (int i = 3; i < 6; ++i) cmat[i + 6 * i] = mfac * 0.5 * (1.0 - 2.0 * nu);
      multiply_nn<6, 6, 1>(S, cmat, glstrain);
      return;
    }
    // ElastHyper: elast_hyper_evaluate, ...service.cpp:19-78
    double C_strain[6], iC_strain[6], prinv[3];
    double dPI[3] = {0, 0, 0}, ddPII[6] = {0, 0, 0, 0, 0, 0};
    for (int i = 0; i < 6; ++i) C_strain[i] = 2.0 * glstrain[i];  // update(2.0, E, 0.0)
    for (int i = 0; i < 3; ++i) C_strain[i] += 1.0;
    voigt_inverse_tensor<true>(C_strain, iC_strain);
    voigt_invariants_principal_strain(prinv, C_strain);
    // elast_hyper_evaluate_invariant_derivatives: loop over the summands (:96-103)
    bool isoprinc = false, anisoprinc = false;
    if (mp.kind == MAT_NEOHOOKE)
    {
      coupneohooke_add_derivatives_principal(mp.young, mp.nue, dPI, ddPII, prinv);
      isoprinc = true;
    }
    else
      for (int k = 0; k < mp.n_summands; ++k)
      {
        if (mp.sum_kind[k] == SUM_COUPNEOHOOKE) coupneohooke_add_derivatives_principal(mp.sum_par[k][0], mp.sum_par[k][1], dPI, ddPII, prinv);
        if (mp.sum_kind[k] == SUM_COUPSVK) coupsvk_add_derivatives_principal(mp.sum_par[k][0], mp.sum_par[k][1], dPI, ddPII, prinv);
        isoprinc = isoprinc || mp.sum_kind[k] != SUM_COUPANISOEXPO;
        anisoprinc = anisoprinc || mp.sum_kind[k] == SUM_COUPANISOEXPO;
      }
    for (int i = 0; i < 6; ++i) S[i] = 0.0;
    for (int i = 0; i < 36; ++i) cmat[i] = 0.0;
    if (isoprinc) elast_hyper_add_isotropic_stress_cmat(S, cmat, C_strain, iC_strain, prinv, dPI, ddPII);
    // elast_hyper_add_anisotropic_princ (:386-396)
    if (anisoprinc)
      for (int k = 0; k < mp.n_summands; ++k)
        if (mp.sum_kind[k] == SUM_COUPANISOEXPO)
          coupanisoexpo_add_stress_aniso_principal(mp.sum_par[k][0], mp.sum_par[k][1], mp.sum_par[k][2], mp.sum_par[k][3], fiber, C_strain, cmat, S);
  }

  // StVenantKirchhoff::strain_energy   stvenantkirchhoff.cpp:183-194 (golden value 908.6538)
  inline double svk_strain_energy(double young, double nue, const double* glstrain)
  {
    MaterialParams mp{MAT_SVK, young, nue};
    double S[6], cmat[36];
    material_evaluate(mp, nullptr, glstrain, S, cmat);
    double psi = 0.0;
    for (int k = 0; k < 6; ++k) psi += glstrain[k] * S[k];
    return psi / 2.0;
  }
}  // namespace fc_oracle
