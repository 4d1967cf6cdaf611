// CPU ORACLE - test infrastructure only (see oracle/README.md).  Restatement of the solid-shell formulations of the reference:
// assumed natural strains (TECH shell_ans) on hex8 and wedge6, optionally with the EAS modes of the solid shell (TECH
// shell_eas_ans: eas_sosh8 with 7 modes on hex8, eas_sosw6 with 1 mode on wedge6).  The element's t direction is the thickness.
//   sampling points, sampling point data     src/solid_3D_ele/4C_solid_3D_ele_calc_shell_ans_lib.hpp:24-47, 506-560
//   local B-operator, local GL strains       ...calc_shell_ans_lib.hpp:61-145, 147-330
//   geometric stiffness                      ...calc_shell_ans_lib.hpp:332-504
//   ShellANSFormulation                      src/solid_3D_ele/4C_solid_3D_ele_calc_shell_ans.hpp:52-196
//   ShellEASANSFormulation                   src/solid_3D_ele/4C_solid_3D_ele_calc_shell_eas_ans.hpp:48-345
//   T^-T at the Gauss point / at the centroid, Mtilde, condensation: as in fc_oracle_eas.hpp
#pragma once
#include "fc_oracle_eas.hpp"

namespace fc_oracle
{
  // history of an EAS-ANS element: alpha_inc [n] | alpha [n] | s [n] | invKaa [n x n] | Kda [nd x n] | old_step_length
  inline int shell_history_doubles(int neas, int nd) { return neas ? 3 * neas + neas * neas + nd * neas + 1 : 0; }
  inline int shell_eas_modes(int ct, int form) { return form == FORM_SHELL_EAS_ANS ? (ct == HEX8 ? 7 : 1) : 0; }

  // Voigt transformation T(J) of the Jacobian J (col-major J[d + 3 i] = dX_i / dxi_d), then its inverse
  inline void voigt_transformation_inverse(const double* Jm, double* TinvT)
  {
    auto J = [&](int i, int j) { return Jm[i + 3 * j]; };
    const int va[6] = {0, 1, 2, 0, 1, 0}, vb[6] = {0, 1, 2, 1, 2, 2};
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
        TinvT[r + 6 * c] = v;
      }
    invert_dense(6, TinvT);
  }

  template <int NEN>
  struct ShellSamplingPoint
  {
    double dN[3 * NEN];   // derivatives at the sampling point, dN[d + 3 a]
    double Jref[9], Jcur[9];  // [d + 3 i]
    double bop[3 * 3 * NEN];  // rows tt, st, rt of the local B-operator: bop[row + 3 (3 a + i)]
  };

  // fe [3 nen], Ke [3 nen x 3 nen] col-major (either may be null); gp_out as in element_evaluate_eas; hist may be null without EAS
  template <int NEN, int NGP>
  inline int element_evaluate_shell_t(int ct, int neas, const MaterialParams& mp, const double* X, const double* u, double* hist,
      double* fe, double* Ke, const double* fiber, int fiber_gp_stride, double* gp_out, int out_kind, int out_type)
  {
    constexpr int ND = 3 * NEN;
    constexpr int NSP = NEN == 8 ? 8 : 5;
    static const double sp_hex[8][3] = {{0, -1, 0}, {1, 0, 0}, {0, 1, 0}, {-1, 0, 0}, {-1, -1, 0}, {1, -1, 0}, {1, 1, 0}, {-1, 1, 0}};
    static const double sp_wedge[5][3] = {{0.5, 0, 0}, {0, 0.5, 0}, {0, 0, 0}, {1, 0, 0}, {0, 1, 0}};
    bool singular = false;
    double* alpha_inc = hist;
    double* alpha = hist ? alpha_inc + neas : nullptr;
    double* svec = hist ? alpha + neas : nullptr;
    double* invKaa = hist ? svec + neas : nullptr;
    double* Kda = hist ? invKaa + neas * neas : nullptr;
    const bool integrate = gp_out == nullptr;
    if (integrate && neas)
    {
      for (int i = 0; i < neas * neas; ++i) invKaa[i] = 0.0;
      for (int i = 0; i < ND * neas; ++i) Kda[i] = 0.0;
      for (int i = 0; i < neas; ++i) svec[i] = 0.0;
    }
    for (int a = 0; a < NEN; ++a)
    {
      double xi[3], dN[3 * NEN];
      node_param_coords(ct, a, xi);
      shape_function_deriv1(ct, xi, dN);
      JacobianMapping<NEN> jm;
      evaluate_jacobian_mapping<NEN>(dN, X, jm, &singular);
      if (!(jm.det > 0)) return ELE_ERR_NEG_JAC_AT_NODE;
    }
    // sampling point data
    ShellSamplingPoint<NEN> sp[NSP];
    for (int k = 0; k < NSP; ++k)
    {
      const double* xi = NEN == 8 ? sp_hex[k] : sp_wedge[k];
      shape_function_deriv1(ct, xi, sp[k].dN);
      multiply_nt<3, NEN, 3>(sp[k].Jref, sp[k].dN, X);
      for (int i = 0; i < 9; ++i) sp[k].Jcur[i] = sp[k].Jref[i];
      multiply_nt_scaled<3, NEN, 3>(1.0, sp[k].Jcur, sp[k].dN, u, 1.0);
      for (int a = 0; a < NEN; ++a)
        for (int i = 0; i < 3; ++i)
        {
          const double* dn = sp[k].dN + 3 * a;
          const double* j = sp[k].Jcur;
          sp[k].bop[0 + 3 * (3 * a + i)] = dn[2] * j[2 + 3 * i];
          sp[k].bop[1 + 3 * (3 * a + i)] = dn[1] * j[2 + 3 * i] + dn[2] * j[1 + 3 * i];
          sp[k].bop[2 + 3 * (3 * a + i)] = dn[0] * j[2 + 3 * i] + dn[2] * j[0 + 3 * i];
        }
    }
    auto metric_diff = [&](int k, int d1, int d2)  // g_d1d2 - G_d1d2 at sampling point k
    {
      double g = 0.0, G = 0.0;
      for (int i = 0; i < 3; ++i)
      {
        g += sp[k].Jcur[d1 + 3 * i] * sp[k].Jcur[d2 + 3 * i];
        G += sp[k].Jref[d1 + 3 * i] * sp[k].Jref[d2 + 3 * i];
      }
      return g - G;
    };
    // EAS: centroid transformation
    double T0invT[36], detJ0 = 0.0;
    if (neas)
    {
      double xi0[3], dN0[3 * NEN];
      centroid_param(ct, xi0);
      shape_function_deriv1(ct, xi0, dN0);
      JacobianMapping<NEN> jm0;
      evaluate_jacobian_mapping<NEN>(dN0, X, jm0, nullptr);
      detJ0 = jm0.det;
      voigt_transformation_inverse(jm0.J, T0invT);
    }
    double gxi[NGP][3], gw[NGP];
    gauss_rule(ct, gxi, gw);
    std::vector<double> M(6 * (neas ? neas : 1)), Mt(6 * (neas ? neas : 1)), cM(6 * (neas ? neas : 1));
    double Bl[6 * ND], B[6 * ND], cb[6 * ND];
    for (int gp = 0; gp < NGP; ++gp)
    {
      const double r = gxi[gp][0], s = gxi[gp][1];
      double dN[3 * NEN];
      shape_function_deriv1(ct, gxi[gp], dN);
      JacobianMapping<NEN> jm;
      evaluate_jacobian_mapping<NEN>(dN, X, jm, &singular);
      const double fac = jm.det * gw[gp];
      double jc[9];
      for (int i = 0; i < 9; ++i) jc[i] = jm.J[i];
      multiply_nt_scaled<3, NEN, 3>(1.0, jc, dN, u, 1.0);
      // interpolation weights of the sampling points for the rows tt (2), st (4), rt (5)
      double wtt[NSP] = {0}, wst[NSP] = {0}, wrt[NSP] = {0};
      if (NEN == 8)
      {
        wtt[4] = 0.25 * (1 - r) * (1 - s);
        wtt[5] = 0.25 * (1 + r) * (1 - s);
        wtt[6] = 0.25 * (1 + r) * (1 + s);
        wtt[7] = 0.25 * (1 - r) * (1 + s);
        wst[1] = 0.5 * (1 + r);
        wst[3] = 0.5 * (1 - r);
        wrt[0] = 0.5 * (1 - s);
        wrt[2] = 0.5 * (1 + s);
      }
      else
      {
        wtt[2] = 1 - r - s;
        wtt[3] = r;
        wtt[4] = s;
        wst[1] = r;
        wrt[0] = s;
      }
      for (int a = 0; a < NEN; ++a)
        for (int i = 0; i < 3; ++i)
        {
          const int c = 3 * a + i;
          const double* dn = dN + 3 * a;
          Bl[0 + 6 * c] = dn[0] * jc[0 + 3 * i];
          Bl[1 + 6 * c] = dn[1] * jc[1 + 3 * i];
          Bl[3 + 6 * c] = dn[0] * jc[1 + 3 * i] + dn[1] * jc[0 + 3 * i];
          double b2 = 0.0, b4 = 0.0, b5 = 0.0;
          for (int k = 0; k < NSP; ++k)
          {
            b2 += wtt[k] * sp[k].bop[0 + 3 * c];
            b4 += wst[k] * sp[k].bop[1 + 3 * c];
            b5 += wrt[k] * sp[k].bop[2 + 3 * c];
          }
          Bl[2 + 6 * c] = b2;
          Bl[4 + 6 * c] = b4;
          Bl[5 + 6 * c] = b5;
        }
      double El[6];
      {
        double g00 = 0, G00 = 0, g11 = 0, G11 = 0, g01 = 0, G01 = 0;
        for (int i = 0; i < 3; ++i)
        {
          g00 += jc[0 + 3 * i] * jc[0 + 3 * i];
          G00 += jm.J[0 + 3 * i] * jm.J[0 + 3 * i];
          g11 += jc[1 + 3 * i] * jc[1 + 3 * i];
          G11 += jm.J[1 + 3 * i] * jm.J[1 + 3 * i];
          g01 += jc[0 + 3 * i] * jc[1 + 3 * i];
          G01 += jm.J[0 + 3 * i] * jm.J[1 + 3 * i];
        }
        El[0] = 0.5 * (g00 - G00);
        El[1] = 0.5 * (g11 - G11);
        El[3] = g01 - G01;
        double e2 = 0.0, e4 = 0.0, e5 = 0.0;
        for (int k = 0; k < NSP; ++k)
        {
          e2 += wtt[k] * metric_diff(k, 2, 2);
          e4 += wst[k] * metric_diff(k, 1, 2);
          e5 += wrt[k] * metric_diff(k, 0, 2);
        }
        El[2] = 0.5 * e2;
        El[4] = e4;
        El[5] = e5;
      }
      double TinvT[36];
      voigt_transformation_inverse(jm.J, TinvT);
      multiply_nn<6, 6, ND>(B, TinvT, Bl);
      double E[6];
      multiply_nn<6, 6, 1>(E, TinvT, El);
      if (neas)
      {
        for (int i = 0; i < 6 * neas; ++i) M[i] = 0.0;
        const double t = gxi[gp][2];
        if (NEN == 8)
          eas_shape_functions(7, gxi[gp], M.data());
        else
          M[2] = t;  // eas_sosw6: one mode, t on the thickness strain
        const double scale = detJ0 / jm.det;
        for (int k = 0; k < neas; ++k)
          for (int rr = 0; rr < 6; ++rr)
          {
            double v = T0invT[rr] * M[6 * k];
            for (int c = 1; c < 6; ++c) v += T0invT[rr + 6 * c] * M[c + 6 * k];
            Mt[rr + 6 * k] = scale * v;
          }
        for (int rr = 0; rr < 6; ++rr)
        {
          double v = Mt[rr] * alpha[0];
          for (int k = 1; k < neas; ++k) v += Mt[rr + 6 * k] * alpha[k];
          E[rr] += v;
        }
      }
      SpatialMaterialMapping<NEN> sm;
      double x[ND];
      for (int i = 0; i < ND; ++i) x[i] = X[i] + u[i];
      evaluate_spatial_material_mapping<NEN>(ct, jm, x, u, 1.0, sm, &singular);
      double Fc[9];
      eas_deformation_gradient(sm.F, E, Fc);
      double S[6], cmat[36];
      material_evaluate(mp, Fc, E, S, cmat, fiber ? fiber + (size_t)fiber_gp_stride * gp : nullptr);
      if (gp_out)
      {
        if (out_kind == 1)
          strain_output(out_type, E, Fc, gp_out + 6 * gp);
        else if (out_type == STRESS_CAUCHY)
          pk2_to_cauchy(S, Fc, gp_out + 6 * gp);
        else
          for (int v = 0; v < 6; ++v) gp_out[6 * gp + v] = S[v];
        continue;
      }
      if (neas)
      {
        for (int k = 0; k < neas; ++k)
          for (int rr = 0; rr < 6; ++rr)
          {
            double v = cmat[rr] * Mt[6 * k];
            for (int c = 1; c < 6; ++c) v += cmat[rr + 6 * c] * Mt[c + 6 * k];
            cM[rr + 6 * k] = v;
          }
        for (int l = 0; l < neas; ++l)
        {
          for (int k = 0; k < neas; ++k)
          {
            double v = 0.0;
            for (int rr = 0; rr < 6; ++rr) v += Mt[rr + 6 * k] * cM[rr + 6 * l];
            invKaa[k + neas * l] += fac * v;
          }
          for (int d = 0; d < ND; ++d)
          {
            double v = 0.0;
            for (int rr = 0; rr < 6; ++rr) v += B[rr + 6 * d] * cM[rr + 6 * l];
            Kda[d + ND * l] += fac * v;
          }
        }
        for (int k = 0; k < neas; ++k)
        {
          double v = 0.0;
          for (int rr = 0; rr < 6; ++rr) v += Mt[rr + 6 * k] * S[rr];
          svec[k] += fac * v;
        }
      }
      if (fe) multiply_tn_scaled<ND, 6, 1>(fac, fe, B, S, 1.0);
      if (Ke)
      {
        multiply_nn<6, 6, ND>(cb, cmat, B);
        multiply_tn_scaled<ND, 6, ND>(fac, Ke, B, cb, 1.0);
        // add_ans_geometric_stiffness: G_ab in natural coordinates, to global with T^-T, contracted with S
        double Sl[6];  // T^-T^T S, so that S . (T^-T G) = Sl . G
        for (int c = 0; c < 6; ++c)
        {
          double v = 0.0;
          for (int rr = 0; rr < 6; ++rr) v += TinvT[rr + 6 * c] * S[rr];
          Sl[c] = v;
        }
        for (int a = 0; a < NEN; ++a)
          for (int b = 0; b < NEN; ++b)
          {
            const double* da = dN + 3 * a;
            const double* db = dN + 3 * b;
            double G[6];
            G[0] = da[0] * db[0];
            G[1] = da[1] * db[1];
            G[3] = da[0] * db[1] + da[1] * db[0];
            double g2 = 0.0, g4 = 0.0, g5 = 0.0;
            for (int k = 0; k < NSP; ++k)
            {
              const double* sa = sp[k].dN + 3 * a;
              const double* sb = sp[k].dN + 3 * b;
              g2 += wtt[k] * sa[2] * sb[2];
              g4 += wst[k] * (sa[1] * sb[2] + sa[2] * sb[1]);
              g5 += wrt[k] * (sa[0] * sb[2] + sa[2] * sb[0]);
            }
            G[2] = g2;
            G[4] = g4;
            G[5] = g5;
            double Gij = 0.0;
            for (int c = 0; c < 6; ++c) Gij += Sl[c] * G[c];
            Gij *= fac;
            for (int d = 0; d < 3; ++d) Ke[(3 * a + d) + ND * (3 * b + d)] += Gij;
          }
      }
    }
    if (!integrate || !neas) return singular ? ELE_ERR_SINGULAR : ELE_OK;
    invert_dense(neas, invKaa);
    std::vector<double> mKi((size_t)ND * neas);
    for (int l = 0; l < neas; ++l)
      for (int d = 0; d < ND; ++d)
      {
        double v = 0.0;
        for (int k = 0; k < neas; ++k) v += Kda[d + ND * k] * invKaa[k + neas * l];
        mKi[d + (size_t)ND * l] = -v;
      }
    if (fe)
      for (int d = 0; d < ND; ++d)
      {
        double v = 0.0;
        for (int k = 0; k < neas; ++k) v += mKi[d + (size_t)ND * k] * svec[k];
        fe[d] += v;
      }
    if (Ke)
      for (int c = 0; c < ND; ++c)
        for (int d = 0; d < ND; ++d)
        {
          double v = 0.0;
          for (int k = 0; k < neas; ++k) v += mKi[d + (size_t)ND * k] * Kda[c + ND * k];
          Ke[d + ND * c] += v;
        }
    return singular ? ELE_ERR_SINGULAR : ELE_OK;
  }

  inline int element_evaluate_shell(int ct, int neas, const MaterialParams& mp, const double* X, const double* u, double* hist, double* fe,
      double* Ke, const double* fiber, int fiber_gp_stride, double* gp_out = nullptr, int out_kind = 0, int out_type = 0)
  {
    if (ct == HEX8) return element_evaluate_shell_t<8, 8>(ct, neas, mp, X, u, hist, fe, Ke, fiber, fiber_gp_stride, gp_out, out_kind, out_type);
    if (ct == WEDGE6) return element_evaluate_shell_t<6, 6>(ct, neas, mp, X, u, hist, fe, Ke, fiber, fiber_gp_stride, gp_out, out_kind, out_type);
    return -1;
  }

  // recover of an EAS-ANS element (update_alpha): alpha_inc = -invKaa (s + Kda^T du); alpha += step_length alpha_inc
  inline void shell_recover(int neas, int nd, double* hist, const double* du, double step_length)
  {
    double* alpha_inc = hist;
    double* alpha = alpha_inc + neas;
    double* svec = alpha + neas;
    double* invKaa = svec + neas;
    double* Kda = invKaa + neas * neas;
    std::vector<double> tmp(neas);
    for (int k = 0; k < neas; ++k)
    {
      double v = svec[k];
      for (int d = 0; d < nd; ++d) v += Kda[d + nd * k] * du[d];
      tmp[k] = v;
    }
    for (int k = 0; k < neas; ++k)
    {
      double v = 0.0;
      for (int l = 0; l < neas; ++l) v += invKaa[k + neas * l] * tmp[l];
      alpha_inc[k] = -v;
    }
    for (int k = 0; k < neas; ++k) alpha[k] += step_length * alpha_inc[k];
    Kda[nd * neas] = step_length;
  }
}  // namespace fc_oracle
