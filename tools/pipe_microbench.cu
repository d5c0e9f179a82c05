// Measures integer-pipe issue rates on the device it runs on: POPC, LOP3, IMAD, VIMNMX and the
// kernel-1 instruction mixes, in thread-operations per clock per SM.  The POPC figure is the
// denominator of kernel 1's roofline (SURVEY.md §8(d) asks to confirm the 16/clk/SM planning
// number on the box).  One CTA of 1024 threads per SM, cycles from clock64() inside the kernel,
// so the result does not depend on the SM clock DVFS picks.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pipe_microbench pipe_microbench.cu
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>

constexpr int ITERS = 4096;
constexpr int CHAINS = 8;

template <int OP>
__global__ void __launch_bounds__(1024) bench(uint32_t seed, unsigned long long* cycles, uint32_t* sink) {
  uint32_t v[CHAINS], w[CHAINS];
#pragma unroll
  for (int i = 0; i < CHAINS; ++i) {
    v[i] = seed * (threadIdx.x + 1) + i * 0x9E3779B9u;
    w[i] = v[i] ^ 0x5bd1e995u;
  }
  __syncthreads();
  const long long t0 = clock64();
#pragma unroll 1
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int i = 0; i < CHAINS; ++i) {
      if (OP == 0) {  // POPC (result fed back so the chain cannot be folded)
        // popc followed by an xor with a live value: the xor issues on the (4x wider) alu pipe,
        // so the chain is paced by POPC and cannot be folded away
        uint32_t p;
        asm volatile("popc.b32 %0, %1;" : "=r"(p) : "r"(v[i]));
        v[i] = p ^ w[i];
        w[i] += seed;
      } else if (OP == 1) {  // LOP3
        asm volatile("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(v[i]) : "r"(v[i]), "r"(w[i]), "r"(seed));
      } else if (OP == 2) {  // IMAD
        asm volatile("mad.lo.u32 %0, %1, %2, %3;" : "=r"(v[i]) : "r"(v[i]), "r"(w[i]), "r"(seed));
      } else if (OP == 3) {  // VIMNMX
        asm volatile("min.u32 %0, %1, %2;" : "=r"(v[i]) : "r"(v[i]), "r"(w[i]));
        asm volatile("max.u32 %0, %1, %2;" : "=r"(w[i]) : "r"(v[i]), "r"(seed));
      } else if (OP == 4) {  // 1 POPC : 2 LOP3 (can both pipes issue together?)
        uint32_t p;
        asm volatile("popc.b32 %0, %1;" : "=r"(p) : "r"(v[i]));
        asm volatile("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(v[i]) : "r"(v[i]), "r"(w[i]), "r"(seed));
        asm volatile("lop3.b32 %0, %1, %2, %3, 0xE8;" : "=r"(w[i]) : "r"(w[i]), "r"(p), "r"(seed));
      } else if (OP == 5) {  // 1 POPC : 4 LOP3 : 1 IMAD : 1 VIMNMX (kernel-1 CSA mix per 1/4 compare)
        uint32_t p, a, b;
        asm volatile("popc.b32 %0, %1;" : "=r"(p) : "r"(v[i]));
        asm volatile("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(a) : "r"(v[i]), "r"(w[i]), "r"(seed));
        asm volatile("lop3.b32 %0, %1, %2, %3, 0xE8;" : "=r"(b) : "r"(v[i]), "r"(w[i]), "r"(seed));
        asm volatile("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(a) : "r"(a), "r"(b), "r"(p));
        asm volatile("lop3.b32 %0, %1, %2, %3, 0xE8;" : "=r"(b) : "r"(a), "r"(b), "r"(p));
        asm volatile("mad.lo.u32 %0, %1, %2, %3;" : "=r"(v[i]) : "r"(p), "r"(seed), "r"(a));
        asm volatile("min.u32 %0, %1, %2;" : "=r"(w[i]) : "r"(w[i]), "r"(b));
      }
    }
  }
  const long long t1 = clock64();
  uint32_t acc = 0;
#pragma unroll
  for (int i = 0; i < CHAINS; ++i) acc ^= v[i] ^ w[i];
  if (acc == 0x12345678u) sink[0] = acc;
  __syncthreads();
  if (threadIdx.x == 0) cycles[blockIdx.x] = (unsigned long long)(t1 - t0);
}

// Floating-point pipes that bound kernel 3: DFMA, the unfused DMUL+DADD pair the parity mode
// issues (numpy rounds the square and the sum separately), and FFMA for the fp32 mode.
template <int OP>
__global__ void __launch_bounds__(1024) benchf(double seed, unsigned long long* cycles, double* sink) {
  double v[CHAINS], w[CHAINS];
  float fv[CHAINS], fw[CHAINS];
#pragma unroll
  for (int i = 0; i < CHAINS; ++i) {
    v[i] = seed * (threadIdx.x + 1) + i * 0.37;
    w[i] = v[i] * 0.99991;
    fv[i] = (float)v[i];
    fw[i] = (float)w[i];
  }
  __syncthreads();
  const long long t0 = clock64();
#pragma unroll 1
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int i = 0; i < CHAINS; ++i) {
      if (OP == 0) {
        asm volatile("fma.rn.f64 %0, %1, %2, %3;" : "=d"(v[i]) : "d"(v[i]), "d"(w[i]), "d"(seed));
      } else if (OP == 1) {
        double d;
        asm volatile("sub.rn.f64 %0, %1, %2;" : "=d"(d) : "d"(v[i]), "d"(w[i]));
        asm volatile("mul.rn.f64 %0, %1, %1;" : "=d"(d) : "d"(d));
        asm volatile("add.rn.f64 %0, %1, %2;" : "=d"(v[i]) : "d"(v[i]), "d"(d));
      } else {
        asm volatile("fma.rn.f32 %0, %1, %2, %3;" : "=f"(fv[i]) : "f"(fv[i]), "f"(fw[i]), "f"((float)seed));
      }
    }
  }
  const long long t1 = clock64();
  double acc = 0;
#pragma unroll
  for (int i = 0; i < CHAINS; ++i) acc += v[i] + w[i] + fv[i] + fw[i];
  if (acc == 0.123456789) sink[0] = acc;
  __syncthreads();
  if (threadIdx.x == 0) cycles[blockIdx.x] = (unsigned long long)(t1 - t0);
}

template <int OP>
double runf(int sms, int ops_per_chain_step, const char* name) {
  unsigned long long* d_cycles;
  double* d_sink;
  cudaMalloc(&d_cycles, sms * sizeof(unsigned long long));
  cudaMalloc(&d_sink, 8);
  benchf<OP><<<sms, 1024>>>(1.000001, d_cycles, d_sink);
  benchf<OP><<<sms, 1024>>>(1.000001, d_cycles, d_sink);
  cudaDeviceSynchronize();
  unsigned long long* h = new unsigned long long[sms];
  cudaMemcpy(h, d_cycles, sms * sizeof(unsigned long long), cudaMemcpyDeviceToHost);
  double mean = 0;
  for (int i = 0; i < sms; ++i) mean += (double)h[i];
  mean /= sms;
  const double rate = 1024.0 * ITERS * CHAINS * ops_per_chain_step / mean;
  printf("  \"%s\": %.2f,\n", name, rate);
  delete[] h;
  cudaFree(d_cycles);
  cudaFree(d_sink);
  return rate;
}

template <int OP>
double run(int sms, int ops_per_chain_step, const char* name) {
  unsigned long long* d_cycles;
  uint32_t* d_sink;
  cudaMalloc(&d_cycles, sms * sizeof(unsigned long long));
  cudaMalloc(&d_sink, 4);
  bench<OP><<<sms, 1024>>>(12345u, d_cycles, d_sink);
  bench<OP><<<sms, 1024>>>(12345u, d_cycles, d_sink);
  cudaDeviceSynchronize();
  unsigned long long* h = new unsigned long long[sms];
  cudaMemcpy(h, d_cycles, sms * sizeof(unsigned long long), cudaMemcpyDeviceToHost);
  double mean = 0;
  for (int i = 0; i < sms; ++i) mean += (double)h[i];
  mean /= sms;
  const double ops = 1024.0 * ITERS * CHAINS * ops_per_chain_step;
  const double rate = ops / mean;
  printf("  \"%s\": %.2f,\n", name, rate);
  delete[] h;
  cudaFree(d_cycles);
  cudaFree(d_sink);
  return rate;
}

int main() {
  cudaDeviceProp p;
  if (cudaGetDeviceProperties(&p, 0) != cudaSuccess) {
    printf("{\"error\": \"no CUDA device\"}\n");
    return 1;
  }
  int khz = 0;
  cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
  printf("{\n  \"device\": \"%s\",\n  \"sm_count\": %d,\n  \"sm_clock_max_khz\": %d,\n", p.name,
         p.multiProcessorCount, khz);
  printf("  \"unit\": \"thread-ops per clock per SM\",\n");
  const int s = p.multiProcessorCount;
  run<0>(s, 1, "popc");
  run<1>(s, 1, "lop3");
  run<2>(s, 1, "imad");
  run<3>(s, 2, "vimnmx");
  run<4>(s, 3, "mix_1popc_2lop3_total");
  run<5>(s, 7, "mix_kernel1_csa_total");
  runf<0>(s, 1, "dfma");
  runf<1>(s, 3, "dsub_dmul_dadd_total");
  runf<2>(s, 1, "ffma");
  printf("  \"note\": \"mix_* count every instruction of the mix\"\n}\n");
  return 0;
}
