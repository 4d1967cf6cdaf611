// C++ host mirror of the reference's call sequence for the path, on top of the C ABI
// (include/fourc_b200.h).  Names and argument meaning follow the reference:
//   Core::FE::Discretization::set_state / evaluate    src/core/fem/src/discretization/4C_fem_discretization.cpp:503-558,
//                                                     4C_fem_discretization_evaluate.cpp:31-103
//   Core::LinAlg::SparseMatrix::zero / complete        src/core/linalg/src/sparse/4C_linalg_sparsematrix.cpp:353-378,843-865
//   Core::Elements::ActionType struct_calc_*           src/solid_3D_ele/4C_solid_3D_ele_evaluate.cpp:82-134
// Errors become exceptions like FOUR_C_THROW (Core::Exception); there is no CPU fallback.
#pragma once
#include <cstdint>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/fourc_b200.h"

namespace FourC::B200
{
  struct Exception : std::runtime_error
  {
    int code;
    Exception(int c, const std::string& m) : std::runtime_error(m), code(c) {}
  };

  enum class CellType { hex8 = FC_HEX8, hex27 = FC_HEX27, tet10 = FC_TET10 };
  enum class ElementTechnology { none = FC_FORM_DISPLACEMENT, fbar = FC_FORM_FBAR };
  enum class ActionType { struct_calc_nlnstiff = FC_ACTION_NLNSTIFF, struct_calc_internalforce = FC_ACTION_INTERNALFORCE };

  // MAT_Struct_StVenantKirchhoff / MAT_ElastHyper + ELAST_CoupNeoHooke (YOUNG, NUE)
  struct Material
  {
    int kind;
    double young, nue;
    static Material StVenantKirchhoff(double young, double nue) { return {FC_MAT_STVENANTKIRCHHOFF, young, nue}; }
    static Material ElastHyperCoupNeoHooke(double young, double nue) { return {FC_MAT_ELASTHYPER_COUPNEOHOOKE, young, nue}; }
  };

  // CSR values of the rank's rows on the device; the graph lives in the Discretization's context.
  class SparseMatrix
  {
   public:
    double* values = nullptr;  // device pointer [nnz], caller-owned (e.g. the Epetra_CrsMatrix values mirrored in HBM)
    int64_t nnz = 0;
  };

  class Discretization
  {
   public:
    Discretization(int device, int myrank, int numproc)
    {
      check(fc_solid_create(&ctx_, device, myrank, numproc), nullptr);
    }
    ~Discretization() { fc_solid_destroy(ctx_); }
    Discretization(const Discretization&) = delete;
    Discretization& operator=(const Discretization&) = delete;

    // column nodes (owned first) + column elements of this rank, as after fill_complete()
    void set_mesh(CellType ct, ElementTechnology tech, const std::vector<double>& X, const std::vector<int>& node_gid,
        const std::vector<int>& node_owner, const std::vector<int>& conn, const std::vector<int>& ele_gid)
    {
      check(fc_solid_set_mesh(ctx_, (int)ct, (int)tech, (int)node_gid.size(), X.data(), node_gid.data(), node_owner.data(),
                (int)ele_gid.size(), conn.data(), ele_gid.data()),
          ctx_);
    }
    void set_material(const Material& m)
    {
      const double p[2] = {m.young, m.nue};
      check(fc_solid_set_material(ctx_, m.kind, p, 2), ctx_);
    }
    // graph of the filled matrix (ExtractMyRowView order) or the full element-connectivity pattern
    void set_graph(int n_row_dof, const int64_t* rowptr, const int* colind, int n_col_dof)
    {
      check(fc_solid_set_graph(ctx_, n_row_dof, rowptr, colind, n_col_dof), ctx_);
    }
    int64_t build_graph()
    {
      check(fc_solid_build_graph(ctx_), ctx_);
      int64_t nnz = 0;
      fc_solid_graph_sizes(ctx_, nullptr, nullptr, &nnz);
      return nnz;
    }
    void set_dirichlet_rows(const std::vector<int>& row_lids)
    {
      check(fc_solid_set_dirichlet_rows(ctx_, (int)row_lids.size(), row_lids.data()), ctx_);
    }

    // set_state(0, "displacement", disnp): the column-layout vector (after the halo import)
    void set_state(const double* u_col) { u_col_ = u_col; }
    void clear_state() { u_col_ = nullptr; }

    // evaluate(params, systemmatrix1, nullptr, systemvector1, nullptr, nullptr): adds into K and f.
    // returns false if elements failed (eval_error_check semantics); throws on a hard error.
    bool evaluate(ActionType action, SparseMatrix* K, double* f_row, int* first_bad_ele_gid = nullptr)
    {
      if (!u_col_) throw Exception(FC_ERR_INVALID, "state 'displacement' not set");
      int bad = 0, gid = -1;
      const int rc = fc_solid_evaluate(ctx_, (int)action, 0, u_col_, f_row, K ? K->values : nullptr, &bad, &gid);
      if (rc == FC_ERR_NEGATIVE_JACOBIAN)
        throw Exception(rc, "determinant of jacobian is <= 0 at one node of element " + std::to_string(gid));
      check(rc, ctx_);
      if (first_bad_ele_gid) *first_bad_ele_gid = gid;
      return bad == 0;
    }
    // SparseMatrix::zero() + Vector::put_scalar(0.0)
    void zero(SparseMatrix* K, double* f_row) { check(fc_solid_zero(ctx_, f_row, K ? K->values : nullptr), ctx_); }

    fc_ctx* handle() { return ctx_; }

   private:
    static void check(int rc, const fc_ctx* c)
    {
      if (rc != FC_OK) throw Exception(rc, fc_last_error(c));
    }
    fc_ctx* ctx_ = nullptr;
    const double* u_col_ = nullptr;
  };
}  // namespace FourC::B200
