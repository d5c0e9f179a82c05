// GPU probe for the tcgen05 path of kernel 1: one 128 x 256 x 256 int8 tile through the same
// expansion kernel, bulk copies, shared-memory descriptors and TMEM loads as hamming_tensor.cu,
// dumped and compared with a host dot product.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I../include -I../opencv-starry-sky-panorama-stitcher_b200/csrc -o build/tc_probe tc_probe.cu
#include "hamming_tensor.cu"
#include <cstdlib>
using namespace ssp;

__global__ void __launch_bounds__(128, 1)
probe_kernel(const uint8_t* a_exp, int a_rows_pad, const uint8_t* b_exp, int b_rows_pad, uint32_t idesc,
             int* __restrict__ out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* sA = smem;
  uint8_t* sB = smem + TC_A_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + TC_A_BYTES + TC_B_BYTES);
  uint32_t* slot = reinterpret_cast<uint32_t*>(bars + 4);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (tid == 0) {
    mbar_init(&bars[0], 1);
    mbar_init(&bars[1], 1);
    fence_mbar_init();
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 256;" ::"r"(smem_u32(slot)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *slot;
  if (tid == 0) {
    mbar_arrive_expect_tx(&bars[0], TC_A_BYTES + TC_B_BYTES);
    tma_bulk_g2s(sA, a_exp, TC_M * 128, &bars[0]);
    tma_bulk_g2s(sA + TC_M * 128, a_exp + (size_t)a_rows_pad * 128, TC_M * 128, &bars[0]);
    tma_bulk_g2s(sB, b_exp, TC_N * 128, &bars[0]);
    tma_bulk_g2s(sB + TC_N * 128, b_exp + (size_t)b_rows_pad * 128, TC_N * 128, &bars[0]);
    tc_wait(&bars[0], 0);
    tc_fence_after();
    const uint32_t a_addr = smem_u32(sA), b_addr = smem_u32(sB);
    for (int k = 0; k < 8; ++k) {
      const uint64_t ad = tc_smem_desc(a_addr + (k >> 2) * (TC_M * 128) + (k & 3) * 32);
      const uint64_t bd = tc_smem_desc(b_addr + (k >> 2) * (TC_N * 128) + (k & 3) * 32);
      tc_mma_i8(tmem, ad, bd, idesc, k > 0 ? 1u : 0u);
    }
    tc_commit(&bars[1]);
  }
  tc_wait(&bars[1], 0);
  tc_fence_after();
  const uint32_t taddr = tmem + ((uint32_t)(warp * 32) << 16);
  for (int c = 0; c < 8; ++c) {
    int v[32];
    tc_ld32(taddr + c * 32, v);
    tc_wait_ld(v);
    for (int j = 0; j < 32; ++j) out[(warp * 32 + lane) * 256 + c * 32 + j] = v[j];
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 256;" ::"r"(tmem) : "memory");
}

int main() {
  const int NA = 128, NB = 256;
  std::vector<uint8_t> a(NA * 32), b(NB * 32);
  srand(1);
  for (auto& x : a) x = rand() & 255;
  for (auto& x : b) x = rand() & 255;
  for (int i = 0; i < 32; ++i) b[5 * 32 + i] = a[3 * 32 + i];  // planted exact match (3, 5)
  uint8_t *d_a, *d_b, *d_ae, *d_be;
  int* d_out;
  cudaMalloc(&d_a, a.size()); cudaMalloc(&d_b, b.size());
  cudaMalloc(&d_ae, 256 * 256); cudaMalloc(&d_be, 256 * 256);
  cudaMalloc(&d_out, 128 * 256 * 4);
  cudaMemcpy(d_a, a.data(), a.size(), cudaMemcpyHostToDevice);
  cudaMemcpy(d_b, b.data(), b.size(), cudaMemcpyHostToDevice);
  expand_bank_kernel<<<(256 * 16 + 255) / 256, 256>>>(d_a, NA, 256, d_ae);
  expand_bank_kernel<<<(256 * 16 + 255) / 256, 256>>>(d_b, NB, 256, d_be);
  // check the expansion itself on the host
  std::vector<uint8_t> ae(256 * 256);
  cudaMemcpy(ae.data(), d_ae, ae.size(), cudaMemcpyDeviceToHost);
  int exp_bad = 0;
  for (int r = 0; r < NA; ++r)
    for (int k = 0; k < 256; ++k) {
      const int bit = (a[r * 32 + k / 8] >> (k % 8)) & 1;
      const int kh = k / 128, kb = k % 128, c8 = kb / 16;
      const size_t off = (size_t)kh * 256 * 128 + (r / 8) * 1024 + (r % 8) * 128 + ((c8 ^ (r % 8)) * 16) + kb % 16;
      const int8_t got = (int8_t)ae[off];
      if (got != (bit ? -1 : 1)) ++exp_bad;
    }
  printf("expansion mismatches: %d\n", exp_bad);
  const size_t smem = TC_A_BYTES + TC_B_BYTES + 256;
  cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  std::vector<int> out(128 * 256);
  const uint32_t variants[2] = {TC_IDESC, TC_IDESC & ~((1u << 7) | (1u << 10))};
  for (int v = 0; v < 2; ++v) {
    cudaMemset(d_out, 0xCD, out.size() * 4);
    probe_kernel<<<1, 128, smem>>>(d_ae, 256, d_be, 256, variants[v], d_out);
    cudaError_t e = cudaDeviceSynchronize();
    printf("variant %d idesc=0x%08x: %s\n", v, variants[v], cudaGetErrorString(e));
    if (e != cudaSuccess) return 1;
    cudaMemcpy(out.data(), d_out, out.size() * 4, cudaMemcpyDeviceToHost);
    int bad = 0;
    for (int m = 0; m < NA; ++m)
      for (int n = 0; n < NB; ++n) {
        int ham = 0;
        for (int i = 0; i < 32; ++i) ham += __builtin_popcount(a[m * 32 + i] ^ b[n * 32 + i]);
        const int want = 256 - 2 * ham;
        if (out[m * 256 + n] != want) {
          if (bad < 6) printf("  D[%d][%d] = %d, want %d\n", m, n, out[m * 256 + n], want);
          ++bad;
        }
      }
    printf("  mismatches: %d of %d; D[3][5] = %d (want 256)\n", bad, NA * NB, out[3 * 256 + 5]);
  }
  return 0;
}
