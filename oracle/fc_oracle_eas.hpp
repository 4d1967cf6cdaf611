// CPU ORACLE - test infrastructure only (see oracle/README.md).  Restatement of the hex8 EAS formulation of the reference.
//   EASFormulation                         src/solid_3D_ele/4C_solid_3D_ele_calc_eas.hpp:44-291
//   EasIterationData / pack order          src/solid_3D_ele/4C_solid_3D_ele_calc_eas_helpers.hpp:91-111, calc_eas.hpp:266-286
//   evaluate_centroid_transformation       ...calc_eas_helpers.hpp:137-155
//   evaluate_voigt_transformation_matrix   src/solid_3D_ele/4C_solid_3D_ele_calc_lib.hpp:1177-1244
//   update_alpha_increment / update_alpha  ...calc_eas_helpers.hpp:189-228
//   EAS shape functions M(xi) (9 / 21)     ...calc_eas_helpers.hpp:258-388
//   Mtilde = detJ0/detJ T0^-T M            ...calc_eas_helpers.hpp:404-420
//   evaluate_gl_strains_with_eas           ...calc_eas_helpers.hpp:478-497
//   integrate_eas                          ...calc_eas_helpers.hpp:511-537
//   condensation                           ...calc_eas.hpp:190-206, calc_eas_helpers.hpp:549-585
//   evaluate_eas_kinematics (nonlinear)    ...calc_eas_helpers.hpp:599-633
//   compute_deformation_gradient_from_gl_strains  ...calc_lib.hpp:1257-1299
//   flow of one evaluation                 src/solid_3D_ele/4C_solid_3D_ele_calc.cpp:110-240; recover :244-273
// Strain output types: src/solid_3D_ele/4C_solid_3D_ele_calc_lib_io.hpp:113-153, 4C_solid_3D_ele_utils.cpp:35-84.
// The reference calls LAPACK (dsyev, dgetrf/dgetri) for the 3x3 eigenproblems and the dense inverses; the oracle uses cyclic
// Jacobi rotations and Gauss-Jordan with partial pivoting - same results to rounding.
#pragma once
#include <cmath>
#include <vector>

#include "fc_oracle_element.hpp"

namespace fc_oracle
{
  // ---- EAS ----------------------------------------------------------------------------------------------------------
  // element history, in the order EASFormulation::pack writes it:
  //   alpha_inc [n] | alpha [n] | s [n] | invKaa [n x n] | Kda [24 x n] | old_step_length      (matrices column-major)
  inline bool eas_linear(int form) { return form == FORM_EAS_MILD_LINEAR || form == FORM_EAS_FULL_LINEAR; }
  inline bool form_is_eas(int form) { return form == FORM_EAS_MILD || form == FORM_EAS_FULL || form == FORM_EAS_SHELL || eas_linear(form); }
  inline int eas_modes(int form) { return (form == FORM_EAS_FULL || form == FORM_EAS_FULL_LINEAR) ? 21 : (form == FORM_EAS_SHELL ? 7 : 9); }
  inline int eas_history_doubles(int neas) { return 3 * neas + neas * neas + 24 * neas + 1; }
  struct EasView
  {
    int n;
    double *alpha_inc, *alpha, *s, *invKaa, *Kda, *old_step;
    EasView(int neas, double* h) : n(neas)
    {
      alpha_inc = h;
      alpha = alpha_inc + n;
      s = alpha + n;
      invKaa = s + n;
      Kda = invKaa + n * n;
      old_step = Kda + 24 * n;
    }
  };
  inline void eas_history_init(int neas, double* h)
  {
    for (int i = 0; i < eas_history_doubles(neas); ++i) h[i] = 0.0;
    h[eas_history_doubles(neas) - 1] = 1.0;
  }

  // M(xi) in parameter space, col-major 6 x n (n = 9: eas_mild, n = 21: eas_full)
  inline void eas_shape_functions(int neas, const double* xi, double* M)
  {
    for (int i = 0; i < 6 * neas; ++i) M[i] = 0.0;
    const double r = xi[0], s = xi[1], t = xi[2];
    auto set = [&](int row, int col, double v) { M[row + 6 * col] = v; };
    if (neas == 7)
    {
      // eas_sosh8: solid shell with t as thickness direction (4C_solid_3D_ele_calc_eas_helpers.hpp:325-349)
      set(0, 0, r); set(1, 1, s); set(2, 2, t);
      set(2, 5, r * t); set(2, 6, s * t);
      set(3, 3, r); set(3, 4, s);
      return;
    }
    // the nine modes both interpolations share
    set(0, 0, r); set(1, 1, s); set(2, 2, t);
    set(3, 3, r); set(3, 4, s);
    set(4, 5, s); set(4, 6, t);
    set(5, 7, r); set(5, 8, t);
    if (neas == 21)
    {
      set(3, 9, r * t);  set(3, 10, s * t);
      set(4, 11, r * s); set(4, 12, r * t);
      set(5, 13, r * s); set(5, 14, s * t);
      set(0, 15, r * s); set(0, 16, r * t);
      set(1, 17, r * s); set(1, 18, s * t);
      set(2, 19, r * t); set(2, 20, s * t);
    }
  }

  // T0^-T (6x6 col-major) and det J0 from the Jacobian at the element centroid
  inline void eas_centroid_transformation(const double* X, double* T0invT, double* detJ0)
  {
    const double xi0[3] = {0.0, 0.0, 0.0};
    double dN[24];
    shape_function_deriv1(HEX8, xi0, dN);
    JacobianMapping<8> jm;
    evaluate_jacobian_mapping<8>(dN, X, jm, nullptr);
    *detJ0 = jm.det;
    auto J = [&](int i, int j) { return jm.J[i + 3 * j]; };
    const int va[6] = {0, 1, 2, 0, 1, 0}, vb[6] = {0, 1, 2, 1, 2, 2};
    // T(r, c) with Voigt pairs r = (a, b), c = (p, q): J(a,p) J(b,q) symmetrised over the row pair; rows of shear strains
    // carry the factor 2 of the strain-like notation
    for (int r = 0; r < 6; ++r)
      for (int c = 0; c < 6; ++c)
      {
        const int a = va[r], b = vb[r], p = va[c], q = vb[c];
        double v;
        if (c < 3)
          v = (r < 3 ? 1.0 : 2.0) * J(a, p) * J(b, p);
        else if (r < 3)
          v = J(a, p) * J(a, q);
        else
          v = J(a, p) * J(b, q) + J(b, p) * J(a, q);
        T0invT[r + 6 * c] = v;
      }
    invert_dense(6, T0invT);
  }

  // F^enh = R U^enh with R from the displacement-based F and U^enh = sqrt(2 E^enh + 1)
  inline void eas_deformation_gradient(const double* F_disp, const double* E_enh, double* F_enh)
  {
    const double Cm[9] = {2.0 * E_enh[0] + 1.0, E_enh[3], E_enh[5], E_enh[3], 2.0 * E_enh[1] + 1.0, E_enh[4], E_enh[5], E_enh[4],
        2.0 * E_enh[2] + 1.0};
    double U_enh[9], Cd[9], U_disp[9], iU[9], R[9];
    sym_function3(Cm, [](double x) { return std::sqrt(x); }, U_enh);
    multiply_tn<3, 3, 3>(Cd, F_disp, F_disp);
    sym_function3(Cd, [](double x) { return std::sqrt(x); }, U_disp);
    invert3x3(iU, U_disp);
    multiply_nn<3, 3, 3>(R, F_disp, iU);
    multiply_nn<3, 3, 3>(F_enh, R, U_enh);
  }

  // recover (struct_calc_recover, default step): alpha_inc = -invKaa (s + Kda^T du); alpha += step_length alpha_inc
  inline void eas_recover(int neas, double* hist, const double* du /* [24] */, double step_length)
  {
    EasView h(neas, hist);
    std::vector<double> tmp(neas);
    for (int k = 0; k < neas; ++k)
    {
      double v = h.s[k];
      for (int d = 0; d < 24; ++d) v += h.Kda[d + 24 * k] * du[d];
      tmp[k] = v;
    }
    for (int k = 0; k < neas; ++k)
    {
      double v = 0.0;
      for (int l = 0; l < neas; ++l) v += h.invKaa[k + neas * l] * tmp[l];
      h.alpha_inc[k] = -v;
    }
    for (int k = 0; k < neas; ++k) h.alpha[k] += step_length * h.alpha_inc[k];
    *h.old_step = step_length;
  }

  // one hex8 EAS element: fe [24], Ke [24 x 24] col-major (either may be null), Gauss-point output gp_out [8][6] (null = none):
  // out_kind 0 = stress of stress_type, 1 = strain of strain type stress_type.  Writes invKaa, Kda, s of the history.
  inline int element_evaluate_eas(int neas, const MaterialParams& mp, const double* X, const double* u, double* hist, double* fe,
      double* Ke, const double* fiber, int fiber_gp_stride, double* gp_out = nullptr, int out_kind = 0, int out_type = 0,
      bool linear = false /* KINEM linear: B at F = I, strain = B u + Mtilde alpha, deformation gradient = I */)
  {
    constexpr int NEN = 8, ND = 24;
    EasView h(neas, hist);
    double x[ND];
    for (int i = 0; i < ND; ++i) x[i] = X[i] + u[i];
    bool singular = false;
    double T0invT[36], detJ0;
    eas_centroid_transformation(X, T0invT, &detJ0);
    // a stress/strain evaluation leaves the iteration data alone (calculate_stress does not integrate the EAS matrices)
    const bool integrate = gp_out == nullptr;
    if (integrate)
    {
      for (int i = 0; i < neas * neas; ++i) h.invKaa[i] = 0.0;
      for (int i = 0; i < 24 * neas; ++i) h.Kda[i] = 0.0;
      for (int i = 0; i < neas; ++i) h.s[i] = 0.0;
    }
    for (int a = 0; a < NEN; ++a)
    {
      double xi[3], dN[3 * NEN];
      node_param_coords(HEX8, a, xi);
      shape_function_deriv1(HEX8, xi, dN);
      JacobianMapping<NEN> jm;
      evaluate_jacobian_mapping<NEN>(dN, X, jm, &singular);
      if (!(jm.det > 0)) return ELE_ERR_NEG_JAC_AT_NODE;
    }
    double gxi[8][3], gw[8];
    gauss_rule(HEX8, gxi, gw);
    std::vector<double> M(6 * neas), Mt(6 * neas), cM(6 * neas);
    double B[6 * ND], cb[6 * ND];
    for (int gp = 0; gp < 8; ++gp)
    {
      double dN[3 * NEN];
      shape_function_deriv1(HEX8, gxi[gp], dN);
      JacobianMapping<NEN> jm;
      evaluate_jacobian_mapping<NEN>(dN, X, jm, &singular);
      const double fac = jm.det * gw[gp];
      SpatialMaterialMapping<NEN> sm;
      evaluate_spatial_material_mapping<NEN>(HEX8, jm, x, u, 1.0, sm, &singular);
      if (linear)
        evaluate_linear_strain_gradient<NEN>(jm.N_XYZ, B);
      else
        evaluate_strain_gradient<NEN>(jm.N_XYZ, sm.F, B);
      eas_shape_functions(neas, gxi[gp], M.data());
      const double scale = detJ0 / jm.det;
      for (int k = 0; k < neas; ++k)
        for (int r = 0; r < 6; ++r)
        {
          double v = T0invT[r] * M[6 * k];
          for (int c = 1; c < 6; ++c) v += T0invT[r + 6 * c] * M[c + 6 * k];
          Mt[r + 6 * k] = scale * v;
        }
      double C[9], E[6], Fenh[9];
      if (linear)
        multiply_nn<6, ND, 1>(E, B, u);  // evaluate_linear_gl_strain
      else
      {
        multiply_tn<3, 3, 3>(C, sm.F, sm.F);
        evaluate_green_lagrange_strain(C, E);
      }
      for (int r = 0; r < 6; ++r)
      {
        double v = Mt[r] * h.alpha[0];
        for (int k = 1; k < neas; ++k) v += Mt[r + 6 * k] * h.alpha[k];
        E[r] += v;
      }
      if (linear)
        for (int i = 0; i < 9; ++i) Fenh[i] = (i % 4 == 0) ? 1.0 : 0.0;
      else
        eas_deformation_gradient(sm.F, E, Fenh);
      double S[6], cmat[36];
      material_evaluate(mp, Fenh, E, S, cmat, fiber ? fiber + (size_t)fiber_gp_stride * gp : nullptr);
      if (gp_out)
      {
        if (out_kind == 1)
          strain_output(out_type, E, Fenh, gp_out + 6 * gp);
        else if (out_type == STRESS_CAUCHY)
          pk2_to_cauchy(S, Fenh, gp_out + 6 * gp);
        else
          for (int v = 0; v < 6; ++v) gp_out[6 * gp + v] = S[v];
        continue;
      }
      // integrate_eas
      for (int k = 0; k < neas; ++k)
        for (int r = 0; r < 6; ++r)
        {
          double v = cmat[r] * Mt[6 * k];
          for (int c = 1; c < 6; ++c) v += cmat[r + 6 * c] * Mt[c + 6 * k];
          cM[r + 6 * k] = v;
        }
      for (int l = 0; l < neas; ++l)
      {
        for (int k = 0; k < neas; ++k)
        {
          double v = 0.0;
          for (int r = 0; r < 6; ++r) v += Mt[r + 6 * k] * cM[r + 6 * l];
          h.invKaa[k + neas * l] += fac * v;
        }
        for (int d = 0; d < ND; ++d)
        {
          double v = 0.0;
          for (int r = 0; r < 6; ++r) v += B[r + 6 * d] * cM[r + 6 * l];
          h.Kda[d + 24 * l] += fac * v;
        }
      }
      for (int k = 0; k < neas; ++k)
      {
        double v = 0.0;
        for (int r = 0; r < 6; ++r) v += Mt[r + 6 * k] * S[r];
        h.s[k] += fac * v;
      }
      if (fe) multiply_tn_scaled<ND, 6, 1>(fac, fe, B, S, 1.0);
      if (Ke)
      {
        multiply_nn<6, 6, ND>(cb, cmat, B);
        multiply_tn_scaled<ND, 6, ND>(fac, Ke, B, cb, 1.0);
        add_geometric_stiffness_matrix<NEN>(jm.N_XYZ, S, fac, Ke);
      }
    }
    if (!integrate) return singular ? ELE_ERR_SINGULAR : ELE_OK;
    // condensation: invKaa <- Kaa^-1, f += -Kda invKaa s, K += -Kda invKaa Kda^T
    invert_dense(neas, h.invKaa);
    std::vector<double> mKi((size_t)ND * neas);
    for (int l = 0; l < neas; ++l)
      for (int d = 0; d < ND; ++d)
      {
        double v = 0.0;
        for (int k = 0; k < neas; ++k) v += h.Kda[d + 24 * k] * h.invKaa[k + neas * l];
        mKi[d + (size_t)ND * l] = -v;
      }
    if (fe)
      for (int d = 0; d < ND; ++d)
      {
        double v = 0.0;
        for (int k = 0; k < neas; ++k) v += mKi[d + (size_t)ND * k] * h.s[k];
        fe[d] += v;
      }
    if (Ke)
      for (int c = 0; c < ND; ++c)
        for (int d = 0; d < ND; ++d)
        {
          double v = 0.0;
          for (int k = 0; k < neas; ++k) v += mKi[d + (size_t)ND * k] * h.Kda[c + 24 * k];
          Ke[d + ND * c] += v;
        }
    return singular ? ELE_ERR_SINGULAR : ELE_OK;
  }
}  // namespace fc_oracle
