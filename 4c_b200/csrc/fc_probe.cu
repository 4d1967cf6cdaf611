// Measurement helpers: FP64 peak probes for the roofline denominators that MEASURED_PEAKS.json
// does not carry (it holds HBM copy bandwidth and bf16 GEMM only).
//   fc_probe_fp64_fma   dependent-chain-free DFMA loop, all SMs           -> TFLOP/s (2 flop per FMA)
//   fc_probe_fp64_mma   mma.sync.m8n8k4.f64 loop (FP64 tensor cores), all SMs -> TFLOP/s
#include <cuda_runtime.h>

#include "../../include/fourc_b200.h"

namespace
{
  __global__ void k_dfma(double* out, int iters, double seed)
  {
    double a[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) a[i] = seed + i + threadIdx.x * 1e-3;
    const double m = 1.0000001, c = 1e-9;
    for (int it = 0; it < iters; ++it)
    {
#pragma unroll
      for (int i = 0; i < 16; ++i) a[i] = fma(a[i], m, c);
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < 16; ++i) s += a[i];
    if (s == 123.456) out[0] = s;
  }

  __global__ void k_dmma(double* out, int iters, double seed)
  {
    double c[8][2];
#pragma unroll
    for (int i = 0; i < 8; ++i) c[i][0] = c[i][1] = seed * i;
    const double a = 1e-3 * (threadIdx.x & 31), b = 1e-3;
    for (int it = 0; it < iters; ++it)
    {
#pragma unroll
      for (int i = 0; i < 8; ++i)
        asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                     : "+d"(c[i][0]), "+d"(c[i][1])
                     : "d"(a), "d"(b));
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += c[i][0] + c[i][1];
    if (s == 123.456) out[0] = s;
  }

  template <class K>
  int run_probe(K kern, int device, double flop_per_thread_iter, double* tflops)
  {
    if (cudaSetDevice(device) != cudaSuccess) return FC_ERR_NO_DEVICE;
    cudaDeviceProp p;
    if (cudaGetDeviceProperties(&p, device) != cudaSuccess) return FC_ERR_CUDA;
    double* d = nullptr;
    if (cudaMalloc((void**)&d, 8) != cudaSuccess) return FC_ERR_CUDA;
    const int threads = 256, blocks = p.multiProcessorCount * 8, iters = 4096;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    double best = 0;
    for (int rep = 0; rep < 5; ++rep)
    {
      cudaEventRecord(e0);
      kern<<<blocks, threads>>>(d, iters, 1.0);
      cudaEventRecord(e1);
      if (cudaEventSynchronize(e1) != cudaSuccess) return FC_ERR_CUDA;
      float ms = 0;
      cudaEventElapsedTime(&ms, e0, e1);
      const double tf = flop_per_thread_iter * iters * (double)threads * blocks / (ms * 1e-3) / 1e12;
      if (tf > best) best = tf;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(d);
    *tflops = best;
    return FC_OK;
  }
}  // namespace

extern "C"
{
  int fc_probe_fp64_fma(int device, double* tflops)
  {
    if (!tflops) return FC_ERR_INVALID;
    return run_probe(k_dfma, device, 16 * 2.0, tflops);
  }
  int fc_probe_fp64_mma(int device, double* tflops)
  {
    if (!tflops) return FC_ERR_INVALID;
    // one m8n8k4 = 8*8*4 FMA per warp = 8 FMA = 16 flop per thread
    return run_probe(k_dmma, device, 8 * 16.0, tflops);
  }
}
