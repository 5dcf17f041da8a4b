// warp_impl.cuh — launcher template shared by warp_i32.cu / warp_f64.cu.
#pragma once
#include <cstdlib>

#include "host_util.h"
#include "kernels_warp.cuh"

namespace alnr {

template <typename T, int R, int MODE, bool MATS>
static int warp_launch_one(cudaStream_t stream, const WarpArgs<T> &a) {
  const unsigned blocks = (a.cv.count + kWarpsPerBlock - 1) / kWarpsPerBlock;
  if (blocks == 0) return ALNR_OK;
  const size_t elems = a.cv.pwm_len ? 4 * (size_t)a.cv.pwm_len : (size_t)a.V * a.V;
  const size_t smem = elems * sizeof(T) * (a.item_matrices ? kWarpsPerBlock : 1);
  if (smem > 48 * 1024)  // long PWM profiles
    ALNR_CUDA(cudaFuncSetAttribute(fill_warp_kernel<T, R, MODE, MATS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  fill_warp_kernel<T, R, MODE, MATS><<<blocks, kWarpsPerBlock * 32, smem, stream>>>(a);
  ALNR_LAUNCHED();
  if (MODE == 2 && !MATS && a.stateful && a.vlist && !a.list) {
    // incremental speculation: settle the mispredicted slots from their top rows, loop over the rest
    warp_verify_kernel<T><<<(a.cv.count + 127) / 128, 128, 0, stream>>>(a);
    ALNR_LAUNCHED();
    WarpArgs<T> b = a;
    b.list = a.rlist; b.list_count = a.rcount;
    b.vlist = nullptr; b.vcount = nullptr;
    fill_warp_kernel<T, R, MODE, MATS><<<blocks, kWarpsPerBlock * 32, smem, stream>>>(b);
    ALNR_LAUNCHED();
  }
  return ALNR_OK;
}

template <typename T>
static int warp_launch_any(cudaStream_t stream, const WarpArgs<T> &a, bool local, bool mats) {
  // MODE 1 (compile-time static penalty) for del == ext.  ALNR_WARP_INT_MODE2=1 sends INT32
  // static-penalty problems through the MODE 2 code instead (the round-1 routing, kept for A/B runs
  // and tests/test_warp.py::test_int32_static_penalty_both_modes; read per launch).
  constexpr bool kStaticMode = true;
  const bool static_ok = sizeof(T) == 8 || getenv("ALNR_WARP_INT_MODE2") == nullptr;
  const int mode = !local ? 0 : ((a.stateful || !static_ok) ? 2 : 1);
  if (mats) {
    if (a.dl.rows_per_word != 8) return ALNR_INVALID_ARGUMENT;
    return mode == 0 ? warp_launch_one<T, 8, 0, true>(stream, a) : warp_launch_one<T, 8, 2, true>(stream, a);
  }
#define ALNR_WARP_CASE(RR)                                                        \
  case RR:                                                                        \
    if (mode == 0) return warp_launch_one<T, RR, 0, false>(stream, a);            \
    if constexpr (kStaticMode) {                                                  \
      if (mode == 1) return warp_launch_one<T, RR, 1, false>(stream, a);          \
    }                                                                             \
    return warp_launch_one<T, RR, 2, false>(stream, a);
  switch (a.dl.rows_per_word) {
    ALNR_WARP_CASE(2) ALNR_WARP_CASE(3) ALNR_WARP_CASE(4) ALNR_WARP_CASE(5) ALNR_WARP_CASE(6)
    ALNR_WARP_CASE(7) ALNR_WARP_CASE(8) ALNR_WARP_CASE(9) ALNR_WARP_CASE(10) ALNR_WARP_CASE(11)
    ALNR_WARP_CASE(12) ALNR_WARP_CASE(13) ALNR_WARP_CASE(14) ALNR_WARP_CASE(15) ALNR_WARP_CASE(16)
    default: return ALNR_INVALID_ARGUMENT;
  }
#undef ALNR_WARP_CASE
}

}  // namespace alnr
