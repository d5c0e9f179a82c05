// Kernel 1, tensor-core variant (tcgen05 kind::i8 on 0/1-expanded descriptors).
// Placeholder until the tcgen05 path lands: the variant reports itself unavailable so callers
// asking for SSP_HAMMING_TENSOR get a clear error instead of a silent substitute.
#include "common.cuh"

namespace ssp {

cudaError_t launch_hamming_tensor(ssp_ctx* ctx, const std::vector<HammingWork>& work, int k,
                                  std::string* why_not) {
  (void)ctx;
  (void)work;
  (void)k;
  *why_not = "tcgen05 kind::i8 kernel not built into this library yet";
  return cudaSuccess;
}

}  // namespace ssp
