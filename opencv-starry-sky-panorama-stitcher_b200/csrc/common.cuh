// Shared host-side context and device-side PTX helpers of libssp_b200.so (sm_100a only).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include <cstdarg>
#include <cstdio>
#include <string>
#include <vector>

#include <nvtx3/nvToolsExt.h>   // header-only; ranges cost nothing unless a profiler is attached

#include "ssp_b200.h"

namespace ssp {

// NVTX range around one stage of a C-ABI call (upload / expand / kernel 1 / finalize / kernel 2 /
// kernel 3 / fetch), so that an nsys / ncu timeline of a step names its pieces.
struct NvtxRange {
  explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
  ~NvtxRange() { nvtxRangePop(); }
  NvtxRange(const NvtxRange&) = delete;
  NvtxRange& operator=(const NvtxRange&) = delete;
};

// Growable device allocation owned by a context.
struct DevBuf {
  void* p = nullptr;
  size_t cap = 0;
  cudaError_t ensure(size_t bytes) {
    if (bytes <= cap) return cudaSuccess;
    if (p) {
      cudaError_t e = cudaFree(p);
      p = nullptr;
      cap = 0;
      if (e != cudaSuccess) return e;
    }
    // grow geometrically so repeated calls with slowly growing sizes do not thrash
    size_t want = bytes + bytes / 4 + 256;
    cudaError_t e = cudaMalloc(&p, want);
    if (e != cudaSuccess) {
      p = nullptr;
      return e;
    }
    cap = want;
    return cudaSuccess;
  }
  void release() {
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
  }
  template <typename T>
  T* as() const {
    return reinterpret_cast<T*>(p);
  }
};

constexpr int SSP_SLOT_DIST_SHIFT = 20;
constexpr int SSP_SLOT_ROW_MASK = (1 << SSP_SLOT_DIST_SHIFT) - 1;

// One directed Hamming work item: every row of `q` against every row of `t`.
struct HammingWork {
  const uint4* q;    // nq rows x 32 B
  const uint4* t;    // nt rows x 32 B
  int nq;
  int nt;
  long long out_row;  // first row of this item in the best2 scratch
  int qi, ti;         // image indices of q and t in the bank
  // tensor variant only: a SPARSE query set.  slot_src[s] = query row held by slot s of the
  // item's 128-slot query tiles, or -1 (s & 7 == row & 7, so rows keep their place in the
  // 8-row swizzle pattern), with an upper bound of the row's nearest-neighbour distance in
  // the bits from SSP_SLOT_DIST_SHIFT up; q_tiles = number of such tiles.  NULL = all nq rows in order.
  const int* slot_src = nullptr;
  int q_tiles = 0;
  // tensor variant, fused cross-check: per train row of the pair the running nearest query row
  // (hamming << 18 | row, atomicMin) and its 16-bit filter threshold; NULL = row direction only
  uint32_t* col_key = nullptr;
  int16_t* col_thr = nullptr;
};

// One image pair for the finalize kernels.
struct HammingPair {
  long long fwd_row;  // best2 rows of query image vs train image
  long long rev_row;  // best2 rows of train image vs query image (cross-check, two-pass forms), or the
                      // pair's first entry in the column-key array (fused cross-check)
  int nq;
  int nt;
};

struct PolyImage {
  DevBuf feat;      // P x 15 float64 (normalised)
  DevBuf star_idx;  // P x 5 int32
  DevBuf angle;     // P x 5 float64 radians
  DevBuf side;      // P x 5 float64 radians
  DevBuf bright;    // P x 5 float64
  long long count = -1;  // -1: not measured yet
  bool has_raw = false;
  bool pending = false;  // measured without waiting: count / domain flag still sit in the pinned slot
};

struct PolyPair {
  const double* f1;
  const double* f2;
  int p1;
  int p2;
  long long out_row;
};

}  // namespace ssp

struct ssp_ctx {
  int device = 0;
  cudaStream_t stream = nullptr;
  cudaStream_t own_stream = nullptr;
  std::string err;
  int sm_count = 0;
  int sm_clock_khz = 0;
  long long hbm_bytes = 0;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;

  // descriptor bank
  ssp::DevBuf desc_owned;
  const uint8_t* d_desc = nullptr;
  std::vector<long long> row_off;

  // Hamming scratch / results
  ssp::DevBuf h_work, h_pairs, h_best2, h_counts, h_offsets, h_mq, h_mt, h_md;
  ssp::DevBuf h_expanded;  // int8 expansion of the bank for the tensor variant (swizzled tiles)
  ssp::DevBuf h_popc;      // popcount of every bank row (int32), tensor variant
  ssp::DevBuf h_tc_work, h_tc_err;
  ssp::DevBuf h_cc_aux, h_cc_mark, h_cc_slots, h_cc_cls;  // sparse reverse pass of the cross-check
  ssp::DevBuf h_col_key, h_col_thr;                        // fused cross-check: column keys / thresholds
  bool h_expanded_valid = false;
  std::vector<long long> exp_byte_off;  // per image byte offset into h_expanded
  std::vector<int> exp_rows_pad;
  std::vector<long long> h_match_offsets;
  long long h_total_rows = 0;   // forward rows of the last call
  long long h_n_matches = -1;
  int h_launches = 0;
  float h_main_ms = 0.f;
  long long h_comparisons = 0;

  // star bank
  ssp::DevBuf s_xy, s_bright, s_rays;
  std::vector<long long> star_off;
  int img_w = 0, img_h = 0;
  double focal = 0.0;

  // polygon banks and scratch
  std::vector<ssp::PolyImage> poly;
  ssp::DevBuf p_cand, p_status, p_scan, p_tmp_idx, p_tmp_angle, p_tmp_side, p_tmp_bright, p_tmp_feat;
  ssp::DevBuf p_pairs, p_argmin, p_mindist, p_counts, p_offsets, p_m1, p_m2, p_md, p_flops;
  // tensor-core screen of kernel 3 (polygon_tensor.cuh): operand banks, |a|^2 table, F per image, tables,
  // candidate pool (+ sqrt values), per-CTA counts, row keys, flags + stats
  ssp::DevBuf t_bank, t_na2, t_f, t_imgs, t_work, t_na2ptr, t_pool, t_sv, t_keep, t_counts, t_rowkey, t_flags;
  long long t_stats_off = 0;    // statistics word behind the F table
  int t_grid = 0;
  int t_used = 0;               // the last ssp_polygon_match_pairs call ran the tensor-core screen
  int t_fallback_flags = 0;     // non-zero: it was abandoned for the CUDA-core kernels (PT_FLAG_*)
  long long t_candidates = -1;
  double t_worst_coef = -1.0;   // max abs(screen - FP64) / F^2 over the candidates of the last call
  long long p_executed_flops = -1;  // flops the last pruning compare executed (device counter), -1 = not counted
  long long p_executed_flops_f32 = -1;  // of those, the single-precision ones (the conservative first-group filter)
  std::vector<long long> p_match_offsets;
  long long p_total_rows = 0;
  long long p_n_matches = -1;
  // pinned slots {count, n_domain} of images measured with SSP_MEASURE_NO_WAIT (polygon.cu)
  long long* p_slots = nullptr;
  int p_slot_capacity = 0;
  bool p_any_pending = false;
  int p_launches = 0;
  float p_main_ms = 0.f;
  long long p_evaluations = 0;

  // brightness scratch
  ssp::DevBuf b_img, b_xy, b_disc, b_out;

  // candidate generation (row N2)
  ssp::DevBuf c_mask, c_work, c_blocks, c_cand;

  // batched RANSAC (row N3)
  ssp::DevBuf r_pairs, r_kp, r_out;
};

namespace ssp {

inline int fail(ssp_ctx* ctx, int code, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  if (ctx) ctx->err = buf;
  return code;
}

#define SSP_CUDA_TRY(ctx, expr)                                                              \
  do {                                                                                       \
    cudaError_t _e = (expr);                                                                 \
    if (_e != cudaSuccess)                                                                   \
      return ssp::fail((ctx), SSP_ERR_CUDA, "%s failed: %s (%s:%d)", #expr,                  \
                       cudaGetErrorString(_e), __FILE__, __LINE__);                          \
  } while (0)

#define SSP_REQUIRE(ctx, cond, code, ...)                    \
  do {                                                       \
    if (!(cond)) return ssp::fail((ctx), (code), __VA_ARGS__); \
  } while (0)

inline cudaMemcpyKind to_device_kind(int mem_kind) {
  return mem_kind == SSP_MEM_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice;
}
inline cudaMemcpyKind from_device_kind(int mem_kind) {
  return mem_kind == SSP_MEM_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost;
}

// ---- device-side PTX helpers (mbarrier + TMA bulk copy) ------------------------------------
#ifdef __CUDACC__
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count)
               : "memory");
}
__device__ __forceinline__ void fence_mbar_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n"
      " .reg .pred p;\n"
      " mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
      " selp.u32 %0, 1, 0, p;\n"
      "}\n"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  while (!mbar_try_wait(bar, parity)) {
  }
}
// TMA 1-D bulk copy global -> shared, completion signalled on `bar` (complete_tx::bytes).
// dst/src 16-byte aligned, bytes a multiple of 16.
__device__ __forceinline__ void tma_bulk_g2s(void* dst, const void* src, uint32_t bytes,
                                             uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
          "r"(smem_u32(dst)),
      "l"(src), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}
// ---- tcgen05 helpers shared by the tensor-core kernels (hamming_tensor.cu, polygon_tensor.cu) ----
// mbarrier wait.  try_wait carries a suspend-time hint so the polling warp sleeps in hardware
// instead of burning issue slots of the SM sub-partition it shares with two epilogue warps (ncu:
// the clock64-per-iteration loop this replaces was 22 % of all issued instructions).  Watchdog by
// spin count: a protocol bug must surface as a CUDA error, not as a hung GPU.
#ifndef TC_WAIT_HINT_NS
#define TC_WAIT_HINT_NS 20000
#endif
template <uint32_t HINT_NS>
__device__ __forceinline__ bool tc_try_wait_hint(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n"
      " .reg .pred p;\n"
      " mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n"
      " selp.u32 %0, 1, 0, p;\n"
      "}\n"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity), "r"(HINT_NS)
      : "memory");
  return ok != 0;
}
template <uint32_t HINT_NS>
__device__ __forceinline__ void tc_wait_hint(uint64_t* bar, uint32_t parity) {
  uint32_t spins = 0;
  while (!tc_try_wait_hint<HINT_NS>(bar, parity)) {
    if (++spins > 100000000u) __trap();
  }
}
__device__ __forceinline__ void tc_wait(uint64_t* bar, uint32_t parity) { tc_wait_hint<TC_WAIT_HINT_NS>(bar, parity); }

// ---- PTX wrappers ---------------------------------------------------------------------------------
__device__ __forceinline__ void tc_fence_before() {
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_after() {
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tc_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(
                   smem_u32(bar))
               : "memory");
}
// K-major, SWIZZLE_128B shared-memory matrix descriptor (cute::UMMA::SmemDescriptor layout):
// start address >> 4 [0,14), LBO >> 4 [16,30) (unused for swizzled K-major), SBO >> 4 [32,46)
// = 1024 B between 8-row groups, version 1 [46,48), layout type 2 = SWIZZLE_128B [61,64).
__device__ __forceinline__ uint64_t tc_smem_desc(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3FFFFu) >> 4);
  d |= (uint64_t)1 << 16;
  d |= (uint64_t)(1024 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
__device__ __forceinline__ void tc_release(uint64_t* acc_empty, int lane) {
  tc_fence_before();
  __syncwarp();
  if (lane == 0) mbar_arrive(acc_empty);
}
// Exclusive scan of n ints into n+1 int64 offsets, single CTA (n is the number of pairs).
static __global__ void __launch_bounds__(1024) scan_counts_kernel(const int* __restrict__ counts, int n,
                                                           long long* __restrict__ offsets) {
  __shared__ long long warp_tot[32];
  __shared__ long long carry_s;
  if (threadIdx.x == 0) carry_s = 0;
  __syncthreads();
  for (int base = 0; base < n; base += 1024) {
    const int i = base + threadIdx.x;
    long long v = (i < n) ? counts[i] : 0;
    long long incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      long long up = __shfl_up_sync(0xffffffffu, incl, o);
      if ((threadIdx.x & 31) >= o) incl += up;
    }
    if ((threadIdx.x & 31) == 31) warp_tot[threadIdx.x >> 5] = incl;
    __syncthreads();
    if (threadIdx.x < 32) {
      long long t = warp_tot[threadIdx.x];
      long long ti = t;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        long long up = __shfl_up_sync(0xffffffffu, ti, o);
        if (threadIdx.x >= o) ti += up;
      }
      warp_tot[threadIdx.x] = ti - t;  // exclusive prefix of warp totals
    }
    __syncthreads();
    const long long carry = carry_s;
    const long long excl = carry + warp_tot[threadIdx.x >> 5] + incl - v;
    if (i < n) offsets[i] = excl;
    __syncthreads();
    if (threadIdx.x == 1023) carry_s = excl + v;
    __syncthreads();
  }
  if (threadIdx.x == 0) offsets[n] = carry_s;
}

#endif  // __CUDACC__

// implemented in api.cu
int use_device(ssp_ctx* ctx);

}  // namespace ssp
