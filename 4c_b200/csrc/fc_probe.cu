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

// ---- RED.F64 throughput vs address pattern (scatter design input) -----------------------------
// Every warp instruction issues 32 lanes; consecutive `span` lanes hit consecutive doubles, each
// span starts on its own 128-byte line.  mode 0: red.add.f64, 1: plain st (same addresses),
// 2: ld+add+st (non-atomic read-modify-write)
namespace
{
  template <int MODE>
  __global__ void k_red_pattern(double* base, long long n_lines, int span, int iters)
  {
    const int lane = threadIdx.x & 31;
    const long long warp = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
    const int groups = 32 / span;  // spans per instruction (lanes beyond groups*span idle)
    const int g = lane / span, w = lane % span;
    if (g >= groups) return;
    for (int it = 0; it < iters; ++it)
    {
      const long long line = ((warp + (long long)it * nwarps) * groups + g) % n_lines;
      double* p = base + line * 16 + w % 16 + (w / 16) * 16 * 0;
      if (span > 16) p = base + ((line * 2) % n_lines) * 16 + w;  // 32 contiguous doubles = 2 lines
      if (MODE == 0)
        atomicAdd(p, 1.0);
      else if (MODE == 1)
        *p = 1.0;
      else
        *p = *p + 1.0;
    }
  }
}  // namespace

extern "C" int fc_probe_red(int device, int mode, int span, long long region_bytes, int iters, double* gops)
{
  if (!gops || span < 1 || span > 32) return FC_ERR_INVALID;
  if (cudaSetDevice(device) != cudaSuccess) return FC_ERR_NO_DEVICE;
  cudaDeviceProp p;
  cudaGetDeviceProperties(&p, device);
  double* d = nullptr;
  if (cudaMalloc((void**)&d, (size_t)region_bytes + 512) != cudaSuccess) return FC_ERR_CUDA;
  cudaMemset(d, 0, (size_t)region_bytes);
  const long long n_lines = region_bytes / 128;
  const int threads = 256, blocks = p.multiProcessorCount * 8;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  double best = 0;
  for (int rep = 0; rep < 3; ++rep)
  {
    cudaEventRecord(e0);
    if (mode == 0) k_red_pattern<0><<<blocks, threads>>>(d, n_lines, span, iters);
    if (mode == 1) k_red_pattern<1><<<blocks, threads>>>(d, n_lines, span, iters);
    if (mode == 2) k_red_pattern<2><<<blocks, threads>>>(d, n_lines, span, iters);
    cudaEventRecord(e1);
    if (cudaEventSynchronize(e1) != cudaSuccess) return FC_ERR_CUDA;
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    const double ops = (double)(32 / span) * span * iters * (double)threads / 32 * blocks;
    best = best > ops / (ms * 1e-3) / 1e9 ? best : ops / (ms * 1e-3) / 1e9;
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  cudaFree(d);
  *gops = best;
  return FC_OK;
}
