// Probe: what does tcgen05.ld.32x32b.xN.pack::16b return?  Writes column c of lane l as
// (l * 1000 + c) with tcgen05.st, reads back packed, prints lane 1's registers.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__global__ void probe(uint32_t* out) {
  __shared__ uint32_t slot;
  const int lane = threadIdx.x;
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 64;" ::"r"(
      (uint32_t)__cvta_generic_to_shared(&slot)) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  __syncwarp();
  const uint32_t t = slot;
  for (int c = 0; c < 64; ++c) {
    uint32_t v = (uint32_t)(lane * 100 + c - 20);   // some negative values too
    asm volatile("tcgen05.st.sync.aligned.32x32b.x1.b32 [%0], {%1};" ::"r"(t + c), "r"(v) : "memory");
  }
  asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
  uint32_t r[8];
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.pack::16b.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
               : "r"(t) : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
  for (int i = 0; i < 8; ++i) out[lane * 8 + i] = r[i];
  __syncwarp();
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 64;" ::"r"(t) : "memory");
}
int main() {
  uint32_t* d; cudaMalloc(&d, 32 * 8 * 4);
  probe<<<1, 32>>>(d);
  uint32_t h[256]; cudaError_t e = cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
  printf("status %s\n", cudaGetErrorString(e));
  for (int lane = 0; lane < 3; ++lane) {
    printf("lane %d:", lane);
    for (int i = 0; i < 8; ++i) printf(" (%d,%d)", (int)(short)(h[lane * 8 + i] & 0xffff), (int)(short)(h[lane * 8 + i] >> 16));
    printf("\n");
  }
  return 0;
}
