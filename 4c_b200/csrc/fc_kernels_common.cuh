// Kernel argument block and scatter helpers shared by the element kernels.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

#include "fc_material.cuh"

namespace fc
{
  struct EvalArgs
  {
    // mesh
    const double* X;       // [n_col_nodes][3]
    const int* conn;       // [n_ele][nen]
    const int* ele_gid;    // [n_ele]
    const int* ele_perm;   // [n_ele] element visited by launch slot k (L2 sweep order) or nullptr = k
    int n_ele;             // end of the element range of this launch
    int ele_offset;        // first launch slot of this launch (FC_EVAL_BOUNDARY), 0 otherwise
    int n_row_nodes;       // column nodes [0, n_row_nodes) are owned (their DOF rows are assembled)
    // tables
    const double* dNtab;   // [ngp + nen + 1][nen][3]
    const double* gpw;     // [ngp]
    // state
    const double* u;       // [n_col_dof]
    // scatter maps
    const int64_t* node_rowbase;  // [n_col_nodes], -1 = not owned
    const int* node_rowlen;       // [n_col_nodes]
    const unsigned char* slots;   // [n_ele][nen][nen]
    const unsigned char* dbc;     // [n_row_dof] or nullptr
    const double* geom;           // hex8 geometry cache [n_ele][72] (k_hex8_geom_cache) or nullptr
    const double* fiber;          // unit fibre directions FIBER1 (anisotropic summands) or nullptr: [n_ele][3] per element
                                  // (fiber_gp = 0) or [n_ele][ngp][3] per Gauss point (fiber_gp = 3, nodal fibres)
    const double* mulf;           // FORM_MULF: [n_ele][ngp][18] history = inverse Jacobian (9, N_xyz = invJ dN) | deformation
                                  // gradient (9) of the prestressed configuration, both column-major; else nullptr
    double* eas;                  // hex8 EAS: [n_ele][EasCfg::hist_doubles] element history (alpha read; Kaa^-1, Kda, s written)
    int fiber_ele, fiber_gp;      // strides in doubles: element e, Gauss point g -> fiber + e * fiber_ele + g * fiber_gp
    // outputs
    double* f;             // [n_row_dof] or nullptr
    double* K;             // [nnz] or nullptr
    double* fe_out;        // element-local outputs (evaluate_elements) or nullptr
    double* Ke_out;
    int* err;              // [2]
    MatParams mat;
  };

  // fire-and-forget FP64 add (RED.E.ADD.F64)
  __device__ __forceinline__ void red_add(double* p, double v) { atomicAdd(p, v); }

  // predicated form: no branch around the reduction
  __device__ __forceinline__ void red_add_if(bool pred, double* p, double v)
  {
    asm volatile(
        "{\n\t.reg .pred q;\n\tsetp.ne.s32 q, %0, 0;\n\t@q red.global.add.f64 [%1], %2;\n\t}"
        :
        : "r"((int)pred), "l"(p), "d"(v)
        : "memory");
  }

  // three rows of one node block (row stride in bytes), unconditional
  __device__ __forceinline__ void red_add3(double* p0, long long stride_bytes, double v0, double v1, double v2)
  {
    asm volatile(
        "{\n\t.reg .b64 p1, p2;\n\tadd.s64 p1, %0, %1;\n\tadd.s64 p2, p1, %1;\n\t"
        "red.global.add.f64 [%0], %2;\n\tred.global.add.f64 [p1], %3;\n\tred.global.add.f64 [p2], %4;\n\t}"
        :
        : "l"(p0), "l"(stride_bytes), "d"(v0), "d"(v1), "d"(v2)
        : "memory");
  }

  // three rows of one node block (row stride in bytes), one predicate (ptxas emits a branch region per reduction)
  __device__ __forceinline__ void red_add3_if(bool pred, double* p0, long long stride_bytes, double v0, double v1, double v2)
  {
    asm volatile(
        "{\n\t.reg .pred q;\n\t.reg .b64 p1, p2;\n\tsetp.ne.s32 q, %0, 0;\n\tadd.s64 p1, %1, %2;\n\tadd.s64 p2, p1, %2;\n\t"
        "@q red.global.add.f64 [%1], %3;\n\t@q red.global.add.f64 [p1], %4;\n\t@q red.global.add.f64 [p2], %5;\n\t}"
        :
        : "r"((int)pred), "l"(p0), "l"(stride_bytes), "d"(v0), "d"(v1), "d"(v2)
        : "memory");
  }

  __device__ __forceinline__ void report_error(int* err, int gid)
  {
    atomicAdd(&err[0], 1);
    atomicMin(&err[1], gid);
  }
}  // namespace fc
