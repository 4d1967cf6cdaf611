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
    int n_ele;
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

  __device__ __forceinline__ void report_error(int* err, int gid)
  {
    atomicAdd(&err[0], 1);
    atomicMin(&err[1], gid);
  }
}  // namespace fc
