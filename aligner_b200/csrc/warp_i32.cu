// warp_i32.cu — INT32 instantiations of the warp-per-pair fill kernel.
#include "warp_impl.cuh"

namespace alnr {
int warp_launch_i32(cudaStream_t stream, const WarpArgs<int> &a, bool local, bool mats) {
  return warp_launch_any<int>(stream, a, local, mats);
}
}  // namespace alnr
