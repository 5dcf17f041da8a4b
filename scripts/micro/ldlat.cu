// latency of one 16-byte load by a whole warp (same address), several flavours
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
template <int MODE> __device__ __forceinline__ uint4 ld(const uint16_t *p) {
  uint4 g;
  if (MODE == 0) asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(g.x), "=r"(g.y), "=r"(g.z), "=r"(g.w) : "l"(p) : "memory");
  if (MODE == 1) asm volatile("ld.relaxed.gpu.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(g.x), "=r"(g.y), "=r"(g.z), "=r"(g.w) : "l"(p) : "memory");
  if (MODE == 2) asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(g.x), "=r"(g.y), "=r"(g.z), "=r"(g.w) : "l"(p) : "memory");
  if (MODE == 3) asm volatile("ld.global.ca.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(g.x), "=r"(g.y), "=r"(g.z), "=r"(g.w) : "l"(p) : "memory");
  if (MODE == 4) asm volatile("ld.weak.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(g.x), "=r"(g.y), "=r"(g.z), "=r"(g.w) : "l"(p) : "memory");
  if (MODE == 5) asm volatile("ld.global.cv.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(g.x), "=r"(g.y), "=r"(g.z), "=r"(g.w) : "l"(p) : "memory");
  return g;
}
template <int MODE> __global__ void k(const uint16_t *row, int iters, int stride, long long *out, int lane0only) {
  long long tot = 0; uint32_t acc = 0;
  for (int i = 0; i < iters; ++i) {
    long long t0 = clock64();
    uint4 g = make_uint4(0,0,0,0);
    if (!lane0only || threadIdx.x == 0) g = ld<MODE>(row + (size_t)i * stride);
    acc += g.x + g.w;
    asm volatile("" :: "r"(acc) : "memory");
    long long t1 = clock64();
    tot += t1 - t0;
  }
  if (threadIdx.x == 0) { out[0] = tot / iters; out[1] = acc; }
}
int main() {
  uint16_t *d; long long *o, h[2];
  size_t n = 1 << 24;
  cudaMalloc(&d, n * 2); cudaMalloc(&o, 16);
  const char *names[] = {"ld.global.cg", "ld.relaxed.gpu", "ld.volatile", "ld.global.ca", "ld.weak.cg", "ld.global.cv"};
  for (int pass = 0; pass < 2; ++pass) {
    for (int mode = 0; mode < 6; ++mode) {
      for (int l0 = 0; l0 < 2; ++l0) {
        cudaMemset(d, 1, 4 << 20);  // 4 MB region: L2 resident
        cudaDeviceSynchronize();
        int iters = 2000, stride = 8;
        switch (mode) {
          case 0: k<0><<<1, 32>>>(d, iters, stride, o, l0); break;
          case 1: k<1><<<1, 32>>>(d, iters, stride, o, l0); break;
          case 2: k<2><<<1, 32>>>(d, iters, stride, o, l0); break;
          case 3: k<3><<<1, 32>>>(d, iters, stride, o, l0); break;
          case 4: k<4><<<1, 32>>>(d, iters, stride, o, l0); break;
          case 5: k<5><<<1, 32>>>(d, iters, stride, o, l0); break;
        }
        cudaMemcpy(h, o, 16, cudaMemcpyDeviceToHost);
        if (pass) printf("%-16s lane0only=%d: %lld cycles per dependent 16-byte load (stride 16 B)\n", names[mode], l0, h[0]);
      }
    }
  }
  // a cold region (DRAM)
  k<0><<<1, 32>>>(d + (8 << 20), 2000, 2048, o, 0);
  cudaMemcpy(h, o, 16, cudaMemcpyDeviceToHost);
  printf("cold, stride 4 KB: %lld cycles\n", h[0]);
  return 0;
}
