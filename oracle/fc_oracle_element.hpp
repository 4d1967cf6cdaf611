// ORACLE — TEST INFRASTRUCTURE ONLY. Never linked into or called by the product library.
//
// Restatement of SolidEleCalc<celltype,Formulation>::evaluate_nonlinear_force_stiffness_mass
// (src/solid_3D_ele/4C_solid_3D_ele_calc.cpp:110-240) for hex8/hex27/tet10 with the
// displacement-based (4C_solid_3D_ele_calc_displacement_based.hpp:45-72, 228-250) and F-bar
// (4C_solid_3D_ele_calc_fbar.hpp:55-111, 164-190; 4C_solid_3D_ele_calc_lib_fbar.hpp:49-154)
// formulations.  Library functions followed, all in src/solid_3D_ele/4C_solid_3D_ele_calc_lib.hpp:
//   evaluate_element_nodes :152-203            evaluate_jacobian_mapping :435-448
//   ensure_positive_jacobian_determinant_at_element_nodes :475-496
//   evaluate_deformation_gradient :585-602     (hex8: F = x N_XYZ^T, others: F = I + u N_XYZ^T)
//   evaluate_spatial_material_mapping :548-567 evaluate_cauchy_green :655-667
//   evaluate_green_lagrange_strain :639-652    evaluate_strain_gradient :708-760
//   add_internal_force_vector :851-860         add_elastic_stiffness_matrix :872-885
//   add_geometric_stiffness_matrix :898-927    add_mass_matrix :940-958
//   for_each_gauss_point :974-993
//
// Parity status: element-level K_e/f_e values are pinned by NO reference test ("parity
// unpinned" at that granularity); they are pinned at solve level by the reference's
// RESULT DESCRIPTION vectors (tests/golden/*.json, TOLERANCE 1e-12) and cross-checked by finite
// differences (K = df/du) in tests/test_oracle_element.py.
#pragma once
#include <vector>

#include "fc_oracle_dense.hpp"
#include "fc_oracle_fem.hpp"
#include "fc_oracle_material.hpp"
#include "fc_oracle_strain.hpp"

namespace fc_oracle
{
  enum Formulation
  {
    FORM_DISP = 0,
    FORM_FBAR = 1,
    FORM_LINEAR = 2,  // KINEM linear: src/solid_3D_ele/4C_solid_3D_ele_calc_displacement_based_linear_kinematics.hpp:36-155
    FORM_MULF = 3,    // PRESTRESS_TECH mulf: src/solid_3D_ele/4C_solid_3D_ele_calc_mulf.hpp:40-196, ..._calc_lib_mulf.hpp:24-84
    FORM_EAS_MILD = 4,  // TECH eas_mild (hex8, 9 modes) and
    FORM_EAS_FULL = 5,  // TECH eas_full (hex8, 21 modes): src/solid_3D_ele/4C_solid_3D_ele_calc_eas.hpp, fc_oracle_eas.hpp
    FORM_EAS_MILD_LINEAR = 10,  // TECH eas_mild / eas_full with KINEM linear (EASSolidIntegrator<hex8, ..., KinemType::linear>,
    FORM_EAS_FULL_LINEAR = 11,  // 4C_solid_3D_ele_factory.hpp:58-61; evaluate_eas_kinematics, calc_eas_helpers.hpp:617-632)
    FORM_SHELL_ANS = 8,      // TECH shell_ans (hex8, wedge6) and
    FORM_SHELL_EAS_ANS = 9,  // TECH shell_eas_ans (hex8: 7 EAS modes, wedge6: 1): fc_oracle_shell.hpp
    FORM_EAS_SHELL = 7,  // TECH shell_eas (hex8, the 7 modes of eas_sosh8, no ANS): EASSolidIntegrator<hex8, eastype_sh8_7>
    FORM_MULF_FBAR = 6  // TECH fbar PRESTRESS_TECH mulf (hex8, pyramid5): src/solid_3D_ele/4C_solid_3D_ele_calc_mulf_fbar.hpp:33-300;
                        // history [ngp + 1][18], the last record belongs to the centroid
  };

  enum EleError
  {
    ELE_OK = 0,
    ELE_ERR_NEG_JAC_AT_NODE = 1,  // FOUR_C_ASSERT_ALWAYS in calc_lib.hpp:492-494
    ELE_ERR_SINGULAR = 2          // invert3x3 throw, fixedsizematrix.hpp:1319
  };

  template <int NEN>
  struct JacobianMapping
  {
    double det;
    double J[9], iJ[9];
    double N_XYZ[3 * NEN];  // 3 x NEN col-major
  };

  template <int NEN>
  struct SpatialMaterialMapping
  {
    double F[9], iF[9], detF;
  };

  // jacobian_ = deriv * X^T ; inverse ; N_XYZ = J^-1 deriv         calc_lib.hpp:435-448
  template <int NEN>
  inline void evaluate_jacobian_mapping(
      const double* dN, const double* X, JacobianMapping<NEN>& jm, bool* singular)
  {
    multiply_nt<3, NEN, 3>(jm.J, dN, X);
    jm.det = invert3x3(jm.iJ, jm.J, singular);
    multiply_nn<3, 3, NEN>(jm.N_XYZ, jm.iJ, dN);
  }

  // calc_lib.hpp:548-605
  template <int NEN>
  inline void evaluate_spatial_material_mapping(int ct, const JacobianMapping<NEN>& jm,
      const double* x, const double* u, double scale, SpatialMaterialMapping<NEN>& sm,
      bool* singular)
  {
    if (ct == HEX8)
    {
      multiply_nt_scaled<3, NEN, 3>(scale, sm.F, x, jm.N_XYZ, 0.0);
    }
    else
    {
      for (int i = 0; i < 9; ++i) sm.F[i] = 0.0;
      sm.F[0] = sm.F[4] = sm.F[8] = 1.0;
      multiply_nt_scaled<3, NEN, 3>(scale, sm.F, u, jm.N_XYZ, scale);
    }
    invert3x3(sm.iF, sm.F, singular);
    sm.detF = determinant3x3(sm.F);
  }

  // B-operator (6 x 3*NEN col-major)                              calc_lib.hpp:708-760
  template <int NEN>
  inline void evaluate_strain_gradient(const double* N_XYZ, const double* F, double* B)
  {
    auto Fm = [&](int r, int c) { return F[r + 3 * c]; };
    auto Nx = [&](int d, int i) { return N_XYZ[d + 3 * i]; };
    for (int i = 0; i < NEN; ++i)
    {
      for (int d = 0; d < 3; ++d)
        for (int e = 0; e < 3; ++e) B[d + 6 * (3 * i + e)] = Fm(e, d) * Nx(d, i);
      for (int e = 0; e < 3; ++e)
      {
        B[3 + 6 * (3 * i + e)] = Fm(e, 0) * Nx(1, i) + Fm(e, 1) * Nx(0, i);
        B[4 + 6 * (3 * i + e)] = Fm(e, 1) * Nx(2, i) + Fm(e, 2) * Nx(1, i);
        B[5 + 6 * (3 * i + e)] = Fm(e, 2) * Nx(0, i) + Fm(e, 0) * Nx(2, i);
      }
    }
  }

  // linear B-operator: the strain gradient at F = I                calc_lib.hpp:773-799
  template <int NEN>
  inline void evaluate_linear_strain_gradient(const double* N_XYZ, double* Bop)
  {
    auto Nx = [&](int d, int i) { return N_XYZ[d + 3 * i]; };
    for (int i = 0; i < NEN; ++i)
    {
      for (int r = 0; r < 6; ++r)
        for (int e = 0; e < 3; ++e) Bop[r + 6 * (3 * i + e)] = 0.0;
      for (int d = 0; d < 3; ++d) Bop[d + 6 * (3 * i + d)] = Nx(d, i);
      Bop[3 + 6 * (3 * i + 0)] = Nx(1, i);
      Bop[3 + 6 * (3 * i + 1)] = Nx(0, i);
      Bop[4 + 6 * (3 * i + 1)] = Nx(2, i);
      Bop[4 + 6 * (3 * i + 2)] = Nx(1, i);
      Bop[5 + 6 * (3 * i + 0)] = Nx(2, i);
      Bop[5 + 6 * (3 * i + 2)] = Nx(0, i);
    }
  }

  // E (strain-like Voigt) from C                                   calc_lib.hpp:639-652
  inline void evaluate_green_lagrange_strain(const double* C, double* E)
  {
    E[0] = 0.5 * (C[0] - 1.0);
    E[1] = 0.5 * (C[4] - 1.0);
    E[2] = 0.5 * (C[8] - 1.0);
    E[3] = C[0 + 3 * 1];
    E[4] = C[1 + 3 * 2];
    E[5] = C[2 + 3 * 0];
  }

  // kgeo                                                           calc_lib.hpp:898-927
  template <int NEN>
  inline void add_geometric_stiffness_matrix(
      const double* B_L, const double* S, double fac, double* K)
  {
    constexpr int ND = 3 * NEN;
    double SmB_L[3];
    for (int inod = 0; inod < NEN; ++inod)
    {
      SmB_L[0] = S[0] * B_L[0 + 3 * inod] + S[3] * B_L[1 + 3 * inod] + S[5] * B_L[2 + 3 * inod];
      SmB_L[1] = S[3] * B_L[0 + 3 * inod] + S[1] * B_L[1 + 3 * inod] + S[4] * B_L[2 + 3 * inod];
      SmB_L[2] = S[5] * B_L[0 + 3 * inod] + S[4] * B_L[1 + 3 * inod] + S[2] * B_L[2 + 3 * inod];
      for (int jnod = 0; jnod < NEN; ++jnod)
      {
        double bopstrbop = 0.0;
        for (int idim = 0; idim < 3; ++idim) bopstrbop += B_L[idim + 3 * jnod] * SmB_L[idim];
        for (int d = 0; d < 3; ++d) K[(3 * inod + d) + ND * (3 * jnod + d)] += fac * bopstrbop;
      }
    }
  }

  // add_mass_matrix                                                       calc_lib.hpp:940-958
  template <int NEN>
  inline void add_mass_matrix(const double* N, double integration_factor, double density, double* Me)
  {
    constexpr int ND = 3 * NEN;
    for (int inod = 0; inod < NEN; ++inod)
    {
      const double ifactor = N[inod] * integration_factor * density;
      for (int jnod = 0; jnod < NEN; ++jnod)
      {
        const double massfactor = N[jnod] * ifactor;
        for (int d = 0; d < 3; ++d) Me[(3 * inod + d) + ND * (3 * jnod + d)] += massfactor;
      }
    }
  }

  // One element.  X,u: node-major [3*a+d] == 3 x NEN col-major.  fe[ND], Ke[ND*ND] col-major,
  // Me optional (consistent mass, density rho).  Outputs are ADDED to (caller zeroes them, like
  // AssembleStrategy::clear_element_storage, 4C_fem_general_assemblestrategy.cpp:74-111).
  // 2nd Piola-Kirchhoff -> Cauchy, stress-Voigt (xx,yy,zz,xy,yz,zx)      4C_solid_3D_ele_utils.cpp:20-33
  inline void pk2_to_cauchy(const double* S, const double* F, double* cauchy)
  {
    const double Sm[9] = {S[0], S[3], S[5], S[3], S[1], S[4], S[5], S[4], S[2]};  // col-major == row-major (symmetric)
    double FS[9], cm[9];
    multiply_nn<3, 3, 3>(FS, F, Sm);
    const double idet = 1.0 / determinant3x3(F);
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j)
      {
        double v = 0.0;
        for (int k = 0; k < 3; ++k) v += FS[i + 3 * k] * F[j + 3 * k];  // multiply_nt
        cm[i + 3 * j] = idet * v;
      }
    cauchy[0] = cm[0]; cauchy[1] = cm[4]; cauchy[2] = cm[8];
    cauchy[3] = cm[0 + 3 * 1]; cauchy[4] = cm[1 + 3 * 2]; cauchy[5] = cm[0 + 3 * 2];
  }

  enum StressType
  {
    STRESS_2PK = 0,
    STRESS_CAUCHY = 1,
    OUTPUT_STRAIN = 16  // + StrainType: the Gauss-point output is the strain of that type (struct_calc_stress writes both)
  };

  template <int NEN, int NGP>
  inline int element_evaluate_t(int ct, int form, const MaterialParams& mp, const double* X,
      const double* u, double* fe, double* Ke, double* Me, double rho, double* gp_stress = nullptr,
      int stress_type = STRESS_CAUCHY, const double* fiber = nullptr, int fiber_gp_stride = 0,
      const double* mulf_hist = nullptr /* FORM_MULF: [ngp][18] = inverse_jacobian (9, column-major) | deformation_gradient (9) */)
  {
    constexpr int ND = 3 * NEN;
    double x[ND];
    for (int i = 0; i < ND; ++i) x[i] = X[i] + u[i];  // current_coordinates.update(1,disp,1)

    bool singular = false;

    // F-bar preparation: centroid jacobian mapping + centroid F          calc_fbar.hpp:55-63
    JacobianMapping<NEN> jm0;
    SpatialMaterialMapping<NEN> sm0;
    const bool fbar_like = form == FORM_FBAR || form == FORM_MULF_FBAR;
    // F = (I + u N_xyz_hist^T) F_hist with N_xyz_hist = invJ_hist dN (evaluate_mulf_spatial_material_mapping, calc_lib_mulf.hpp:34-84)
    auto mulf_mapping = [&](const double* hist, const double* dN, SpatialMaterialMapping<NEN>& sm)
    {
      double Nh[3 * NEN], delta[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
      multiply_nn<3, 3, NEN>(Nh, hist, dN);
      multiply_nt_scaled<3, NEN, 3>(1.0, delta, u, Nh, 1.0);
      multiply_nn<3, 3, 3>(sm.F, delta, hist + 9);
      invert3x3(sm.iF, sm.F, &singular);
      sm.detF = determinant3x3(sm.F);
    };
    if (fbar_like)
    {
      double xi0[3], dN0[3 * NEN];
      centroid_param(ct, xi0);
      shape_function_deriv1(ct, xi0, dN0);
      evaluate_jacobian_mapping<NEN>(dN0, X, jm0, &singular);
      evaluate_spatial_material_mapping<NEN>(ct, jm0, x, u, 1.0, sm0, &singular);
      // MulfFBarFormulation::prepare (calc_mulf_fbar.hpp:118-143): N_XYZ_0 stays that of the reference Jacobian, the centroid
      // deformation gradient comes from the centroid's history record
      if (form == FORM_MULF_FBAR) mulf_mapping(mulf_hist + 18 * NGP, dN0, sm0);
    }

    // negative Jacobian check at the element nodes                   calc_lib.hpp:475-496
    for (int a = 0; a < NEN; ++a)
    {
      double xi[3], dN[3 * NEN];
      node_param_coords(ct, a, xi);
      shape_function_deriv1(ct, xi, dN);
      JacobianMapping<NEN> jm;
      evaluate_jacobian_mapping<NEN>(dN, X, jm, &singular);
      if (!(jm.det > 0)) return ELE_ERR_NEG_JAC_AT_NODE;
    }

    double gxi[NGP][3], gw[NGP];
    gauss_rule(ct, gxi, gw);
    double element_mass = 0.0, element_volume = 0.0;

    std::vector<double> Bv(6 * ND), cbv(6 * ND);
    double* B = Bv.data();
    double* cb = cbv.data();

    for (int gp = 0; gp < NGP; ++gp)
    {
      double N[NEN], dN[3 * NEN];
      shape_function(ct, gxi[gp], N);
      shape_function_deriv1(ct, gxi[gp], dN);
      JacobianMapping<NEN> jm;
      evaluate_jacobian_mapping<NEN>(dN, X, jm, &singular);
      const double integration_factor = jm.det * gw[gp];

      SpatialMaterialMapping<NEN> sm;
      evaluate_spatial_material_mapping<NEN>(ct, jm, x, u, 1.0, sm, &singular);

      double Fmat[9], C[9], E[6];  // what the material sees
      double fbar_factor = 1.0, Hop[ND], Cg[9];
      if (form == FORM_MULF)
      {
        // MulfFormulation::evaluate (calc_mulf.hpp:52-90): F = (I + u N_xyz_hist^T) F_hist with N_xyz_hist = invJ_hist dN
        // (calc_lib_mulf.hpp:34-84); B-operator and geometric stiffness with the N_XYZ of the reference Jacobian mapping
        const double* invJh = mulf_hist + 18 * gp;
        const double* Fh = invJh + 9;
        double Nh[3 * NEN], delta[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
        multiply_nn<3, 3, NEN>(Nh, invJh, dN);
        multiply_nt_scaled<3, NEN, 3>(1.0, delta, u, Nh, 1.0);
        multiply_nn<3, 3, 3>(sm.F, delta, Fh);
        invert3x3(sm.iF, sm.F, &singular);
        sm.detF = determinant3x3(sm.F);
        for (int i = 0; i < 9; ++i) Fmat[i] = sm.F[i];
        multiply_tn<3, 3, 3>(C, sm.F, sm.F);
        evaluate_green_lagrange_strain(C, E);
        evaluate_strain_gradient<NEN>(jm.N_XYZ, sm.F, B);
      }
      else if (form == FORM_LINEAR)
      {
        // DisplacementBasedLinearKinematicsFormulation::evaluate: gl_strain = B_lin u (calc_lib.hpp:683-695), the material
        // sees the identity as deformation gradient; no geometric stiffness (add_stiffness_matrix, :138-151)
        evaluate_linear_strain_gradient<NEN>(jm.N_XYZ, B);
        for (int i = 0; i < 9; ++i) Fmat[i] = (i % 4 == 0) ? 1.0 : 0.0;
        multiply_nn<6, ND, 1>(E, B, u);
      }
      else if (form == FORM_DISP)
      {
        for (int i = 0; i < 9; ++i) Fmat[i] = sm.F[i];
        multiply_tn<3, 3, 3>(C, sm.F, sm.F);
        evaluate_green_lagrange_strain(C, E);
        evaluate_strain_gradient<NEN>(jm.N_XYZ, sm.F, B);
      }
      else
      {
        if (form == FORM_MULF_FBAR) mulf_mapping(mulf_hist + 18 * gp, dN, sm);  // MulfFBarFormulation::evaluate, calc_mulf_fbar.hpp:162-166
        fbar_factor = std::pow(sm0.detF / sm.detF, 1.0 / 3.0);  // lib_fbar.hpp:49-53
        evaluate_strain_gradient<NEN>(jm.N_XYZ, sm.F, B);
        // H operator                                                 lib_fbar.hpp:71-104
        double iF0[9], iFg[9];
        invert3x3(iF0, sm0.F, &singular);
        invert3x3(iFg, sm.F, &singular);
        for (int idof = 0; idof < ND; ++idof)
        {
          Hop[idof] = 0.0;
          for (int idim = 0; idim < 3; ++idim)
          {
            Hop[idof] += iF0[idim + 3 * (idof % 3)] * jm0.N_XYZ[idim + 3 * (idof / 3)];
            Hop[idof] -= iFg[idim + 3 * (idof % 3)] * jm.N_XYZ[idim + 3 * (idof / 3)];
          }
        }
        multiply_tn<3, 3, 3>(Cg, sm.F, sm.F);  // linearization.cauchygreen (unmodified F)
        SpatialMaterialMapping<NEN> smbar;
        if (form == FORM_MULF_FBAR)
          for (int i = 0; i < 9; ++i) smbar.F[i] = fbar_factor * sm.F[i];  // get_spatial_material_mapping_bar, calc_mulf_fbar.hpp:59-70
        else
          evaluate_spatial_material_mapping<NEN>(ct, jm, x, u, fbar_factor, smbar, &singular);
        for (int i = 0; i < 9; ++i) Fmat[i] = smbar.F[i];
        multiply_tn<3, 3, 3>(C, smbar.F, smbar.F);
        evaluate_green_lagrange_strain(C, E);
      }

      double S[6], cmat[36];
      // fibre of the Gauss point: the element's FIBER1 (stride 0) or the nodal fibres projected to this point (stride 3)
      material_evaluate(mp, Fmat, E, S, cmat, fiber ? fiber + (size_t)fiber_gp_stride * gp : nullptr);
      if (gp_stress)
      {
        // calculate_stress / assemble_stress_type_to_matrix_row: solid_3D_ele_calc.cpp:347-390, calc_lib_io.hpp:167-190
        if (stress_type >= OUTPUT_STRAIN)
          strain_output(stress_type - OUTPUT_STRAIN, E, Fmat, gp_stress + 6 * gp);
        else if (stress_type == STRESS_CAUCHY)
          pk2_to_cauchy(S, Fmat, gp_stress + 6 * gp);
        else
          for (int v = 0; v < 6; ++v) gp_stress[6 * gp + v] = S[v];
      }

      const double fac_force = fbar_like ? integration_factor / fbar_factor : integration_factor;
      // f += fac * B^T S                                               calc_lib.hpp:851-860
      if (fe) multiply_tn_scaled<ND, 6, 1>(fac_force, fe, B, S, 1.0);

      if (Ke)
      {
        const double fac_mat = fbar_like ? integration_factor * fbar_factor : integration_factor;
        const double fac_geo = fbar_like ? integration_factor / fbar_factor : integration_factor;
        multiply_nn<6, 6, ND>(cb, cmat, B);                              // cb = C B
        multiply_tn_scaled<ND, 6, ND>(fac_mat, Ke, B, cb, 1.0);          // K += fac B^T cb
        if (form != FORM_LINEAR) add_geometric_stiffness_matrix<NEN>(jm.N_XYZ, S, fac_geo, Ke);
        if (fbar_like)
        {
          // add_fbar_stiffness_matrix                                   lib_fbar.hpp:120-154
          double rcg[6], ccg[6];
          voigt_matrix_to_vector_strain(Cg, rcg);
          multiply_nn<6, 6, 1>(ccg, cmat, rcg);
          double bopccg[ND], bops[ND];
          multiply_tn_scaled<ND, 6, 1>(integration_factor * fbar_factor / 3.0, bopccg, B, ccg, 0.0);
          multiply_tn_scaled<ND, 6, 1>(-integration_factor / fbar_factor / 3.0, bops, B, S, 0.0);
          for (int idof = 0; idof < ND; ++idof)
            for (int jdof = 0; jdof < ND; ++jdof)
              Ke[idof + ND * jdof] += Hop[jdof] * (bops[idof] + bopccg[idof]);
        }
      }
      if (Me)
      {
        if (ct != TET10 && ct != TET4)
          add_mass_matrix<NEN>(N, integration_factor, rho, Me);  // same rule as the stiffness integral
        else
        {
          element_mass += rho * integration_factor;  // solid_3D_ele_calc.cpp:201-205
          element_volume += integration_factor;
        }
      }
    }
    if (Me && (ct == TET10 || ct == TET4))
    {
      // mass rule differs from the stiffness rule: second loop with the mean density   solid_3D_ele_calc.cpp:228-240
      double mxi[11][3], mw[11];
      const int nm = gauss_rule_mass(ct, mxi, mw);
      for (int gp = 0; gp < nm; ++gp)
      {
        double N[NEN], dN[3 * NEN];
        shape_function(ct, mxi[gp], N);
        shape_function_deriv1(ct, mxi[gp], dN);
        JacobianMapping<NEN> jm;
        evaluate_jacobian_mapping<NEN>(dN, X, jm, &singular);
        add_mass_matrix<NEN>(N, jm.det * mw[gp], element_mass / element_volume, Me);
      }
    }
    return singular ? ELE_ERR_SINGULAR : ELE_OK;
  }

  // Nodal fibres -> Gauss points of the stiffness rule: interpolated with the element's shape functions and normalised
  // (Core::Nodes::project_fibers_to_gauss_points, src/core/fem/src/general/node/4C_fem_general_fiber_node_utils.cpp:18-75;
  // called from interpolate_fibers_to_gauss_points_and_add_to_parameter_list, 4C_solid_3D_ele_calc_lib.hpp:1110-1143).
  // nodal [nen][3] -> gp_fiber [ngp][3]
  inline void project_fibers_to_gauss_points(int ct, const double* nodal, double* gp_fiber)
  {
    const int nen = num_nodes(ct), ngp = num_gp(ct);
    double xi[27][3], w[27];
    gauss_rule(ct, xi, w);
    for (int g = 0; g < ngp; ++g)
    {
      double N[27], v[3] = {0.0, 0.0, 0.0};
      shape_function(ct, xi[g], N);
      for (int a = 0; a < nen; ++a)
        for (int i = 0; i < 3; ++i) v[i] += N[a] * nodal[3 * a + i];
      const double nrm = std::sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
      for (int i = 0; i < 3; ++i) gp_fiber[3 * g + i] = v[i] * (1.0 / nrm);
    }
  }

  // MULF history of one element, [ngp][18]: inverse_jacobian | deformation_gradient, both 3x3 column-major.
  // init: what MulfFormulation::evaluate stores at its first call (calc_mulf.hpp:59-63) and the identity of MulfHistoryData.
  // update: MulfFormulation::update_prestress (calc_mulf.hpp:150-171) with the displacements u of the finished step.
  template <int NEN, int NGP>
  inline void mulf_history_t(int ct, const double* X, const double* u /* nullptr = init */, double* hist, bool centroid)
  {
    double gxi[NGP + 1][3], gw[NGP];
    gauss_rule(ct, gxi, gw);
    centroid_param(ct, gxi[NGP]);  // MULF with F-bar: one more record for the element centroid (calc_mulf_fbar.hpp:118-143, 255-270)
    for (int gp = 0; gp < NGP + (centroid ? 1 : 0); ++gp)
    {
      double dN[3 * NEN];
      shape_function_deriv1(ct, gxi[gp], dN);
      double* invJh = hist + 18 * gp;
      double* Fh = invJh + 9;
      if (!u)
      {
        JacobianMapping<NEN> jm;
        evaluate_jacobian_mapping<NEN>(dN, X, jm, nullptr);
        for (int i = 0; i < 9; ++i)
        {
          invJh[i] = jm.iJ[i];
          Fh[i] = (i % 4 == 0) ? 1.0 : 0.0;
        }
        continue;
      }
      double Nh[3 * NEN], delta[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1}, Ftot[9], idelta[9], invJ_new[9];
      multiply_nn<3, 3, NEN>(Nh, invJh, dN);
      multiply_nt_scaled<3, NEN, 3>(1.0, delta, u, Nh, 1.0);  // evaluate_mulf_deformation_gradient_update
      multiply_nn<3, 3, 3>(Ftot, delta, Fh);                   // the deformation gradient the evaluation saw
      invert3x3(idelta, delta, nullptr);
      multiply_tn<3, 3, 3>(invJ_new, idelta, invJh);
      for (int i = 0; i < 9; ++i)
      {
        Fh[i] = Ftot[i];
        invJh[i] = invJ_new[i];
      }
    }
  }
  // number of history records of one element
  inline int mulf_records(int ct, int form) { return form == FORM_MULF ? num_gp(ct) : (form == FORM_MULF_FBAR ? num_gp(ct) + 1 : 0); }
  inline void mulf_history(int ct, const double* X, const double* u, double* hist, bool centroid = false)
  {
    switch (ct)
    {
      case HEX8: return mulf_history_t<8, 8>(ct, X, u, hist, centroid);
      case HEX27: return mulf_history_t<27, 27>(ct, X, u, hist, centroid);
      case TET10: return mulf_history_t<10, 4>(ct, X, u, hist, centroid);
      case HEX20: return mulf_history_t<20, 27>(ct, X, u, hist, centroid);
      case TET4: return mulf_history_t<4, 1>(ct, X, u, hist, centroid);
      case WEDGE6: return mulf_history_t<6, 6>(ct, X, u, hist, centroid);
      case PYRAMID5: return mulf_history_t<5, 8>(ct, X, u, hist, centroid);
      case HEX18: return mulf_history_t<18, 18>(ct, X, u, hist, centroid);
      case NURBS27: return mulf_history_t<27, 27>(ct, X, u, hist, centroid);
    }
  }

  inline int element_evaluate(int ct, int form, const MaterialParams& mp, const double* X,
      const double* u, double* fe, double* Ke, double* Me = nullptr, double rho = 0.0, double* gp_stress = nullptr,
      int stress_type = STRESS_CAUCHY, const double* fiber = nullptr, int fiber_gp_stride = 0, const double* mulf_hist = nullptr)
  {
    switch (ct)
    {
      case HEX8:
        return element_evaluate_t<8, 8>(ct, form, mp, X, u, fe, Ke, Me, rho, gp_stress, stress_type, fiber, fiber_gp_stride, mulf_hist);
      case HEX27:
        return element_evaluate_t<27, 27>(ct, form, mp, X, u, fe, Ke, Me, rho, gp_stress, stress_type, fiber, fiber_gp_stride, mulf_hist);
      case TET10:
        return element_evaluate_t<10, 4>(ct, form, mp, X, u, fe, Ke, Me, rho, gp_stress, stress_type, fiber, fiber_gp_stride, mulf_hist);
      case HEX20:
        return element_evaluate_t<20, 27>(ct, form, mp, X, u, fe, Ke, Me, rho, gp_stress, stress_type, fiber, fiber_gp_stride, mulf_hist);
      case TET4:
        return element_evaluate_t<4, 1>(ct, form, mp, X, u, fe, Ke, Me, rho, gp_stress, stress_type, fiber, fiber_gp_stride, mulf_hist);
      case WEDGE6:
        return element_evaluate_t<6, 6>(ct, form, mp, X, u, fe, Ke, Me, rho, gp_stress, stress_type, fiber, fiber_gp_stride, mulf_hist);
      case PYRAMID5:
        return element_evaluate_t<5, 8>(ct, form, mp, X, u, fe, Ke, Me, rho, gp_stress, stress_type, fiber, fiber_gp_stride, mulf_hist);
      case HEX18:
        return element_evaluate_t<18, 18>(ct, form, mp, X, u, fe, Ke, Me, rho, gp_stress, stress_type, fiber, fiber_gp_stride, mulf_hist);
      case NURBS27:  // knots and weights of the element: nurbs_context()
        return element_evaluate_t<27, 27>(ct, form, mp, X, u, fe, Ke, Me, rho, gp_stress, stress_type, fiber, fiber_gp_stride, mulf_hist);
    }
    return -1;
  }
}  // namespace fc_oracle
