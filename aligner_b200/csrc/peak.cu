// peak.cu — INT32 issue-rate micro-benchmarks used as the roofline denominator
// (MEASURED_PEAKS.json has no integer peak, so it is measured here).
// "op" = one 32-bit-lane integer instruction (IADD3 / LOP3 / VIMNMX / IMAD).
#include <algorithm>

#include "host_util.h"

namespace alnr {

// alu pipe only: (x + k) ^ m chains -> IADD3 + LOP3
__global__ void __launch_bounds__(256) peak_alu_kernel(int *out, int k, int m, int iters) {
  int x[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) x[i] = threadIdx.x + i;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int u = 0; u < 4; ++u)
#pragma unroll
      for (int i = 0; i < 8; ++i) x[i] = max(x[i] + k, m) ^ k;
  }
  int s = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) s ^= x[i];
  if (s == 0x7fffffff) out[0] = s;
}

// alu + fma pipes: IADD3/LOP3 chains next to IMAD chains, 1:1
__global__ void __launch_bounds__(256) peak_mixed_kernel(int *out, int k, int m, int iters) {
  int x[4], y[8];
#pragma unroll
  for (int i = 0; i < 4; ++i) x[i] = threadIdx.x + i;
#pragma unroll
  for (int i = 0; i < 8; ++i) y[i] = threadIdx.x * 3 + i;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int u = 0; u < 4; ++u) {
#pragma unroll
      for (int i = 0; i < 4; ++i) x[i] = (x[i] + k) ^ m;   // 2 alu-pipe ops
#pragma unroll
      for (int i = 0; i < 8; ++i) y[i] = y[i] * k + m;     // 1 fma-pipe op each
    }
  }
  int s = 0;
#pragma unroll
  for (int i = 0; i < 4; ++i) s ^= x[i];
#pragma unroll
  for (int i = 0; i < 8; ++i) s ^= y[i];
  if (s == 0x7fffffff) out[0] = s;
}

// fp64 pipe: independent DADD chains (add = 1 op), and DADD + fmax chains (the recurrence's mix)
__global__ void __launch_bounds__(256) peak_f64_add_kernel(double *out, double k, int iters) {
  double x[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) x[i] = (double)(threadIdx.x + i);
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int u = 0; u < 4; ++u)
#pragma unroll
      for (int i = 0; i < 8; ++i) x[i] = __dadd_rn(x[i], k);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += x[i];
  if (s == 1.25) out[0] = s;
}
__global__ void __launch_bounds__(256) peak_f64_addmax_kernel(double *out, double k, double m, int iters) {
  double x[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) x[i] = (double)(threadIdx.x + i);
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int u = 0; u < 4; ++u)
#pragma unroll
      for (int i = 0; i < 8; ++i) x[i] = fmax(__dadd_rn(x[i], k), m);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += x[i];
  if (s == 1.25) out[0] = s;
}

}  // namespace alnr

extern "C" int alnr_measure_fp64_peak(int device, double *add_ops_per_s, double *addmax_ops_per_s) {
  using namespace alnr;
  ALNR_CUDA(cudaSetDevice(device < 0 ? 0 : device));
  DevBuf<double> out;
  ALNR_CUDA(out.alloc(4));
  cudaEvent_t e0, e1;
  ALNR_CUDA(cudaEventCreate(&e0));
  ALNR_CUDA(cudaEventCreate(&e1));
  const int blocks = 148 * 8, threads = 256, iters = 512;
  double best_add = 0, best_mix = 0;
  for (int rep = 0; rep < 4; ++rep) {
    float ms = 0.f;
    ALNR_CUDA(cudaEventRecord(e0));
    peak_f64_add_kernel<<<blocks, threads>>>(out.p, 1.0 + rep, iters);
    ALNR_LAUNCHED();
    ALNR_CUDA(cudaEventRecord(e1));
    ALNR_CUDA(cudaEventSynchronize(e1));
    ALNR_CUDA(cudaEventElapsedTime(&ms, e0, e1));
    double ops = (double)blocks * threads * iters * 4.0 * 8.0;
    if (rep) best_add = std::max(best_add, ops / (ms * 1e-3));
    ALNR_CUDA(cudaEventRecord(e0));
    peak_f64_addmax_kernel<<<blocks, threads>>>(out.p, 1.0 + rep, 1e300, iters);
    ALNR_LAUNCHED();
    ALNR_CUDA(cudaEventRecord(e1));
    ALNR_CUDA(cudaEventSynchronize(e1));
    ALNR_CUDA(cudaEventElapsedTime(&ms, e0, e1));
    ops = (double)blocks * threads * iters * 4.0 * 8.0 * 2.0;  // add + max
    if (rep) best_mix = std::max(best_mix, ops / (ms * 1e-3));
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  if (add_ops_per_s) *add_ops_per_s = best_add;
  if (addmax_ops_per_s) *addmax_ops_per_s = best_mix;
  return ALNR_OK;
}

extern "C" int alnr_measure_int_peak(int device, double *alu_ops_per_s, double *mixed_ops_per_s) {
  using namespace alnr;
  ALNR_CUDA(cudaSetDevice(device < 0 ? 0 : device));
  DevBuf<int> out;
  ALNR_CUDA(out.alloc(4));
  cudaEvent_t e0, e1;
  ALNR_CUDA(cudaEventCreate(&e0));
  ALNR_CUDA(cudaEventCreate(&e1));
  const int blocks = 148 * 8, threads = 256, iters = 2048;
  double best_alu = 0, best_mixed = 0;
  for (int rep = 0; rep < 4; ++rep) {
    float ms = 0.f;
    ALNR_CUDA(cudaEventRecord(e0));
    peak_alu_kernel<<<blocks, threads>>>(out.p, 3 + rep, 5, iters);
    ALNR_LAUNCHED();
    ALNR_CUDA(cudaEventRecord(e1));
    ALNR_CUDA(cudaEventSynchronize(e1));
    ALNR_CUDA(cudaEventElapsedTime(&ms, e0, e1));
    double ops = (double)blocks * threads * iters * 4.0 * 8.0 * 3.0;  // add, max, xor
    if (rep) best_alu = std::max(best_alu, ops / (ms * 1e-3));
    ALNR_CUDA(cudaEventRecord(e0));
    peak_mixed_kernel<<<blocks, threads>>>(out.p, 3 + rep, 5, iters);
    ALNR_LAUNCHED();
    ALNR_CUDA(cudaEventRecord(e1));
    ALNR_CUDA(cudaEventSynchronize(e1));
    ALNR_CUDA(cudaEventElapsedTime(&ms, e0, e1));
    ops = (double)blocks * threads * iters * 4.0 * (4.0 * 2.0 + 8.0);
    if (rep) best_mixed = std::max(best_mixed, ops / (ms * 1e-3));
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  if (alu_ops_per_s) *alu_ops_per_s = best_alu;
  if (mixed_ops_per_s) *mixed_ops_per_s = best_mixed;
  return ALNR_OK;
}
