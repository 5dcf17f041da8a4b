// warp_f64.cu — FP64 instantiations of the warp-per-pair fill kernel (heuristic matrix search,
// non-integral problems).  Built with -fmad=false like the rest of the library.
#include "warp_impl.cuh"

namespace alnr {
int warp_launch_f64(cudaStream_t stream, const WarpArgs<double> &a, bool local, bool mats) {
  return warp_launch_any<double>(stream, a, local, mats);
}
}  // namespace alnr
