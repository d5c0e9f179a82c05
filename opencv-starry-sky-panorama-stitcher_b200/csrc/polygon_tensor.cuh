// Kernel 3, tensor-core SCREEN (round 2): which columns can hold a row's nearest polygon is decided by a
// TF32 contraction on the 5th-generation tensor cores; what the reference returns -- argmin and the FP64
// distance of every row (custom_bf_polygon_matcher.py:992-1034, :1063-1067) -- is then computed in FP64, in
// numpy's operations and order, for exactly those columns.  Included by polygon.cu (one translation unit).
//
// Screen.  dist^2(a, b) = |a|^2 + |b|^2 - 2 a.b, and for a fixed row only  D(a, b) = |b|^2 - 2 a.b  matters.
// Every feature x is split into two TF32 numbers x_h + x_l (22 significand bits), |b|^2 likewise, and
//   A row (64 floats) = [ a_h (15) | 1 | 1 | a_h (15) | a_l (15) | 0 ... ]
//   B row (64 floats) = [-2b_h (15)|n_h|n_l|-2b_l (15)|-2b_h (15)| 0 ... ]          n = |b|^2
// so that one K = 64 contraction gives D = n_h + n_l - 2 (a_h.b_h + a_h.b_l + a_l.b_h): everything but the
// a_l.b_l term and the split residuals.  Both forms of every image are built once per call
// (poly_tf32_bank_kernel) in the shared-memory image of the UMMA descriptor (K-major, 128-byte swizzle, two
// K halves: the layout of kernel 1's expanded descriptor bank), so kernel 1's machinery carries over: one
// persistent CTA per SM, a TMA producer warp (row tile once per item, column tiles through a three-stage
// ring), one thread issuing 8 x tcgen05.mma kind::tf32 (M128 N256 K8) per stage into one of two TMEM
// accumulators, and sixteen epilogue warps, each thread owning one row and one 64-column quarter.
//
// Epilogue.  A thread keeps theta = min over the columns it took of (D + 3E), E an upper bound of
// abs(D - D_real) (below).  A column with D >= theta has dist^2 larger than that of a column already
// taken by at least E, far more than FP64 rounding and sqrt ties can bridge: it is neither the row's minimum
// nor a tie of it.  A column with D < theta is a CANDIDATE: (pair, row, column, D) goes to this CTA's
// segment of a global pool.  The four threads of a row exchange theta through shared memory.  Expected
// candidates per row: the records of a random sequence, ~ln(P2) + a few.
//
// Exact part.  poly_cand_filter_kernel + poly_cand_eval_kernel evaluate every candidate whose D lies below the row's FINAL theta (the
// threads leave it in a per-row word at the end of an item; what was taken while theta was still loose and
// lies at or above it is ruled out like a column never taken) with poly_dist2_f64<0> and takes the row
// minimum of sqrt(v) (atomicMin on the bits of a non-negative double); poly_cand_argmin_kernel gives the row
// the LOWEST column among the candidates that reach that minimum -- numpy's argmin; rows without a finite
// candidate keep (0, inf) like the per-lane kernel.
//
// Error bound E = PT_ERR_COEF * F^2, F >= every abs(feature) of both images (computed with the banks).
//   split residuals: abs(x - x_h - x_l) <= 2^-22 abs(x) (TF32 keeps 11 significand bits, round to nearest),
//   so the dropped terms of 2 a.b are below 2 * 15 * 3.1 * 2^-22 F^2 = 2.3e-5 F^2, those of n below 4.5e-6 F^2;
//   tensor-core accumulation: the 22-bit products are exact in FP32; the adder is not IEEE.  Assumed: each of
//   the 8 instructions errs by at most 2^-21 of the magnitudes it adds (<= 90 F^2): 3.4e-4 F^2.  The assumption
//   is CHECKED on every call: the exact part compares |a|^2 + D with the FP64 value of every candidate and
//   the call falls back to the CUDA-core kernels if any differs by more than E (flag 8), as it does for
//   features too large for a float (flag 2) or a pool overflow (flag 4; adversarial column orders).
//   Observed maximum on the bench workloads: see profiles/ (r02_poly_tensor_screen.txt).
#pragma once

namespace ssp {

constexpr int PT_M = 128;                     // rows per tile = TMEM lanes
constexpr int PT_N = 256;                     // columns per stage = TMEM columns per accumulator
constexpr int PT_ACC = 2;
#ifndef PT_WGS
#define PT_WGS 4                              // epilogue warpgroups: each drains PT_N / PT_WGS columns of every accumulator
                                              // (4: 9.4 ms at the cap against 10.05 with 2 -- a warp held up on its candidate path
                                              // delays the hand-back of half as many columns; 23 % more candidates)
#endif
constexpr int PT_HALF = PT_N / PT_WGS;        // columns of an accumulator per epilogue thread
constexpr int PT_EPI_WARPS = 4 * PT_WGS;
constexpr int PT_STAGES = 3;
constexpr int PT_ROW_BYTES = 256;             // 64 floats
constexpr int PT_A_BYTES = PT_M * PT_ROW_BYTES;
constexpr int PT_B_BYTES = PT_N * PT_ROW_BYTES;
constexpr int PT_THREADS = 32 * PT_EPI_WARPS + 64;   // epilogue warps, producer, MMA issuer
constexpr int PT_PAD = 256;                   // rows of every operand bank are padded to this
constexpr int PT_TAIL_BYTES = 256 + PT_WGS * PT_M * (int)sizeof(float) + 16;   // barriers, theta exchange [warpgroup][row], counter
constexpr size_t PT_SMEM_BYTES = (size_t)PT_A_BYTES + (size_t)PT_STAGES * PT_B_BYTES + PT_TAIL_BYTES;
#ifndef PT_WAIT_HINT_NS
#define PT_WAIT_HINT_NS 20000   // suspend-time hint of the mbarrier waits
#endif
__device__ __forceinline__ void pt_wait(uint64_t* bar, uint32_t parity) { tc_wait_hint<PT_WAIT_HINT_NS>(bar, parity); }
#ifndef PT_EXPERIMENT
#define PT_EXPERIMENT 0   // probes (results are wrong): 1 = accumulators handed back unread (MMA + operand feed only), 2 = read but not examined
#endif
constexpr float PT_ERR_COEF = 5e-4f;          // E = PT_ERR_COEF * F^2
constexpr float PT_F_MAX = 1e15f;             // beyond this F^2 leaves the float range: no screen
constexpr int PT_FLAG_SMEM = 1, PT_FLAG_RANGE = 2, PT_FLAG_OVERFLOW = 4, PT_FLAG_ERROR = 8;
// instruction descriptor: c_format F32 = 1 [4,6), a/b format TF32 = 2 [7,10) / [10,13), both K-major,
// N >> 3 [17,23), M >> 4 [24,29)
constexpr uint32_t PT_IDESC = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(PT_N >> 3) << 17) |
                              ((uint32_t)(PT_M >> 4) << 24);

struct PtBankImg {
  const double* feat;   // count x 15
  long long out_off;    // byte offset of the image's A form; the B form follows rows_pad * 256 bytes behind
  long long row0;       // first entry of the image in the |a|^2 table
  int count, rows_pad;
  int slot;             // entry of the image in the F table
  int pad_;
};
struct PtWork {          // one image pair
  const uint8_t* a_exp;
  const uint8_t* b_exp;
  const float* f_a;
  const float* f_b;
  long long out_row;     // first row of the pair in the per-row outputs
  int p1, p2;
  int a_rows_pad, b_rows_pad;
  int tile_begin;        // prefix sum of row tiles over the work list
  int pair;
};
struct __align__(16) PtCand {
  int pair, row, col;
  float d;
};

// order-preserving map float -> unsigned (for atomicMin on values of either sign) and back
__device__ __forceinline__ unsigned pt_ordered(float x) {
  const unsigned b = __float_as_uint(x);
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float pt_unordered(unsigned k) {
  return __uint_as_float((k & 0x80000000u) ? (k & 0x7FFFFFFFu) : ~k);
}

__device__ __forceinline__ uint32_t pt_tf32(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return r;
}

// ---- operand banks ---------------------------------------------------------------------------------
// one thread per (padded row, 16-byte chunk = 4 of the 64 floats) of image blockIdx.y, both forms
__global__ void __launch_bounds__(256)
poly_tf32_bank_kernel(const PtBankImg* __restrict__ imgs, uint8_t* __restrict__ out_base,
                      double* __restrict__ na2_base, unsigned* __restrict__ f_bits) {
  const PtBankImg im = imgs[blockIdx.y];
  const long long gid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const int row = (int)(gid >> 4);
  const int c = (int)(gid & 15);
  if (row >= im.rows_pad) return;
  uint32_t va[4] = {0, 0, 0, 0}, vb[4] = {0, 0, 0, 0};
  if (row < im.count) {
    double x[15];
    double n = 0.0;
#pragma unroll
    for (int k = 0; k < 15; ++k) {
      x[k] = im.feat[(long long)row * 15 + k];
      n = fma(x[k], x[k], n);
    }
    if (c == 0) {
      na2_base[im.row0 + row] = n;
      float f = 0.0f;
#pragma unroll
      for (int k = 0; k < 15; ++k) f = fmaxf(f, __double2float_ru(fabs(x[k])));
      atomicMax(&f_bits[im.slot], __float_as_uint(f));   // non-negative floats order like their bits
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int k = 4 * c + j;
      float a = 0.0f, b = 0.0f;
      if (k == 15 || k == 16) {
        const float nh = __uint_as_float(pt_tf32(__double2float_rn(n)));
        a = 1.0f;
        b = (k == 15) ? nh : __uint_as_float(pt_tf32(__double2float_rn(n - (double)nh)));
      } else if (k < 47) {
        const int f = k < 15 ? k : (k < 32 ? k - 17 : k - 32);
        const double xv = x[f];
        const float xh = __uint_as_float(pt_tf32(__double2float_rn(xv)));
        const float xl = __uint_as_float(pt_tf32(__double2float_rn(xv - (double)xh)));
        a = (k < 32) ? xh : xl;
        b = -2.0f * ((k >= 17 && k < 32) ? xl : xh);
      }
      va[j] = __float_as_uint(a);
      vb[j] = __float_as_uint(b);
    }
  } else if (c == 3) {
    vb[3] = __float_as_uint(3.0e38f);   // k = 15 of a padding column: D = 3e38 against every real row
  }
  const int kh = c >> 3, c8 = c & 7;
  const size_t off = (size_t)kh * im.rows_pad * 128 + (size_t)(row >> 3) * 1024 + (size_t)(row & 7) * 128 +
                     (size_t)((c8 ^ (row & 7)) * 16);
  uint8_t* out = out_base + im.out_off;
  *reinterpret_cast<uint4*>(out + off) = make_uint4(va[0], va[1], va[2], va[3]);
  *reinterpret_cast<uint4*>(out + (size_t)im.rows_pad * PT_ROW_BYTES + off) = make_uint4(vb[0], vb[1], vb[2], vb[3]);
}

// ---- screen kernel -----------------------------------------------------------------------------------
// D[tmem] (+)= A[smem] * B[smem]^T, TF32 x TF32 -> FP32, M128 N256 K8
__device__ __forceinline__ void pt_mma_tf32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                            uint32_t accumulate) {
  asm volatile(
      "{\n"
      " .reg .pred p;\n"
      " setp.ne.b32 p, %4, 0;\n"
      " tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n"
      "}\n" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// 32 consecutive TMEM columns of this thread's lane
template <int BASE>
__device__ __forceinline__ void pt_ld32(uint32_t taddr, uint32_t (&v)[PT_HALF]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
               : "=r"(v[BASE + 0]), "=r"(v[BASE + 1]), "=r"(v[BASE + 2]), "=r"(v[BASE + 3]), "=r"(v[BASE + 4]), "=r"(v[BASE + 5]), "=r"(v[BASE + 6]), "=r"(v[BASE + 7]), "=r"(v[BASE + 8]), "=r"(v[BASE + 9]), "=r"(v[BASE + 10]), "=r"(v[BASE + 11]), "=r"(v[BASE + 12]), "=r"(v[BASE + 13]), "=r"(v[BASE + 14]), "=r"(v[BASE + 15]), "=r"(v[BASE + 16]), "=r"(v[BASE + 17]), "=r"(v[BASE + 18]), "=r"(v[BASE + 19]), "=r"(v[BASE + 20]), "=r"(v[BASE + 21]), "=r"(v[BASE + 22]), "=r"(v[BASE + 23]), "=r"(v[BASE + 24]), "=r"(v[BASE + 25]), "=r"(v[BASE + 26]), "=r"(v[BASE + 27]), "=r"(v[BASE + 28]), "=r"(v[BASE + 29]), "=r"(v[BASE + 30]), "=r"(v[BASE + 31])
               : "r"(taddr + BASE)
               : "memory");
}
template <int BASE>
__device__ __forceinline__ void pt_wait_ld32(uint32_t (&v)[PT_HALF]) {
  asm volatile("tcgen05.wait::ld.sync.aligned;"
               : "+r"(v[BASE + 0]), "+r"(v[BASE + 1]), "+r"(v[BASE + 2]), "+r"(v[BASE + 3]), "+r"(v[BASE + 4]), "+r"(v[BASE + 5]), "+r"(v[BASE + 6]), "+r"(v[BASE + 7]), "+r"(v[BASE + 8]), "+r"(v[BASE + 9]), "+r"(v[BASE + 10]), "+r"(v[BASE + 11]), "+r"(v[BASE + 12]), "+r"(v[BASE + 13]), "+r"(v[BASE + 14]), "+r"(v[BASE + 15]), "+r"(v[BASE + 16]), "+r"(v[BASE + 17]), "+r"(v[BASE + 18]), "+r"(v[BASE + 19]), "+r"(v[BASE + 20]), "+r"(v[BASE + 21]), "+r"(v[BASE + 22]), "+r"(v[BASE + 23]), "+r"(v[BASE + 24]), "+r"(v[BASE + 25]), "+r"(v[BASE + 26]), "+r"(v[BASE + 27]), "+r"(v[BASE + 28]), "+r"(v[BASE + 29]), "+r"(v[BASE + 30]), "+r"(v[BASE + 31])
               :
               : "memory");
}

__device__ __forceinline__ int pt_find_work(const PtWork* __restrict__ work, int n_work, int tile) {
  int lo = 0, hi = n_work - 1;
  while (lo < hi) {
    const int mid = (lo + hi + 1) >> 1;
    if (work[mid].tile_begin <= tile) lo = mid; else hi = mid - 1;
  }
  return lo;
}

// The columns [BASE, BASE + 32) of a thread's half tile against its theta.  Common path: a minimum tree
// (FMNMX3) and one compare.  A lane finds a candidate in ~3 % of its chunks, a warp in every second one, so
// the other path is kept SHORT and free of branches per column (the first version unrolled 128 test-and-push
// blocks: ncu showed 60 % of the stall samples waiting for instruction fetch): a bit mask of the columns
// below theta, then one push per set bit.  Only a chunk with ONE candidate knows that candidate's D (it is
// the minimum); the others carry NaN and are left out of the screen-against-FP64 check.
template <int BASE>
__device__ __forceinline__ void pt_chunk(const uint32_t (&v)[PT_HALF], float& theta, float e3, const PtWork& w,
                                         int row, int col0, unsigned* cand_n, PtCand* __restrict__ seg,
                                         unsigned seg_cap, int* __restrict__ flags, volatile float* my_slot) {
  float m = __uint_as_float(v[BASE]);
#pragma unroll
  for (int i = 1; i < 32; ++i) m = fminf(m, __uint_as_float(v[BASE + i]));
  const int last_col = w.p2 - (col0 + BASE);          // columns at or beyond this offset are padding (D = 3e38)
  if (m < theta && last_col > 0) {
    // the chunk's minimum is taken (it is below theta and, being a real column's, below every padding
    // column), so what is not within 3E of IT is already ruled out: compare with the new theta at once
    theta = fminf(theta, m + e3);
    *my_slot = theta;
    unsigned mask = 0u;
#pragma unroll
    for (int i = 0; i < 32; ++i) mask |= (__uint_as_float(v[BASE + i]) < theta) ? (1u << i) : 0u;
    if (last_col < 32) mask &= (1u << last_col) - 1u;
    const float d = __popc(mask) == 1 ? m : __int_as_float(0x7fc00000);
    while (mask) {
      const int i = __ffs((int)mask) - 1;
      mask &= mask - 1u;
      const unsigned idx = atomicAdd(cand_n, 1u);
      if (idx < seg_cap) {
        PtCand cnd;
        cnd.pair = w.pair;
        cnd.row = row;
        cnd.col = col0 + BASE + i;
        cnd.d = d;
        seg[idx] = cnd;
      } else {
        atomicOr(flags, PT_FLAG_OVERFLOW);
      }
    }
  }
}

__global__ void __launch_bounds__(PT_THREADS, 1)
poly_screen_tcgen05_kernel(const PtWork* __restrict__ work, int n_work, int total_tiles,
                           PtCand* __restrict__ pool, unsigned seg_cap, unsigned* __restrict__ cand_counts,
                           unsigned* __restrict__ row_theta, int* __restrict__ flags) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* sA = smem;
  uint8_t* sB = smem + PT_A_BYTES;
  uint8_t* tail = smem + PT_A_BYTES + PT_STAGES * PT_B_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(tail);
  uint64_t* a_full = bars + 0;
  uint64_t* a_empty = bars + 1;
  uint64_t* b_full = bars + 2;
  uint64_t* b_empty = b_full + PT_STAGES;
  uint64_t* acc_full = b_empty + PT_STAGES;
  uint64_t* acc_empty = acc_full + PT_ACC;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + PT_ACC);
  static_assert((2 + 2 * PT_STAGES + 2 * PT_ACC) * 8 + 4 <= 256, "barrier block overflows its 256 bytes");
  float* share = reinterpret_cast<float*>(tail + 256);                                    // [PT_WGS][128]
  unsigned* cand_n = reinterpret_cast<unsigned*>(tail + 256 + PT_WGS * PT_M * sizeof(float));

  const int tid = threadIdx.x;
  const int warp = tid >> 5;
  const int lane = tid & 31;

  if ((smem_u32(smem) & 1023u) != 0) {  // SWIZZLE_128B needs 1024-byte aligned tiles
    if (tid == 0) {
      atomicOr(flags, PT_FLAG_SMEM);
      cand_counts[blockIdx.x] = 0;
    }
    return;
  }
  if (tid == 0) {
    mbar_init(a_full, 1);
    mbar_init(a_empty, 1);
    for (int s = 0; s < PT_STAGES; ++s) {
      mbar_init(&b_full[s], 1);
      mbar_init(&b_empty[s], 1);
    }
    for (int b = 0; b < PT_ACC; ++b) {
      mbar_init(&acc_full[b], 1);
      mbar_init(&acc_empty[b], PT_EPI_WARPS);  // one arrive per epilogue warp
    }
    fence_mbar_init();
    *cand_n = 0u;
  }
  for (int i = tid; i < PT_WGS * PT_M; i += PT_THREADS) share[i] = INFINITY;
  if (warp == PT_EPI_WARPS + 1) {  // TMEM: all 512 columns (two accumulators of 256 FP32 columns)
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(tmem_slot))
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  PtCand* seg = pool + (size_t)blockIdx.x * seg_cap;

  if (warp == PT_EPI_WARPS) {
    // ===== TMA producer =====
    if (lane == 0) {
      uint32_t g = 0, item = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++item) {
        const PtWork w = work[pt_find_work(work, n_work, tile)];
        const int r_tile = tile - w.tile_begin;
        if (item > 0) pt_wait(a_empty, (item - 1) & 1u);
        mbar_arrive_expect_tx(a_full, PT_A_BYTES);
        tma_bulk_g2s(sA, w.a_exp + (size_t)r_tile * (PT_M * 128), PT_M * 128, a_full);
        tma_bulk_g2s(sA + PT_M * 128, w.a_exp + (size_t)w.a_rows_pad * 128 + (size_t)r_tile * (PT_M * 128),
                     PT_M * 128, a_full);
        const int n_t = (w.p2 + PT_N - 1) / PT_N;
        for (int t = 0; t < n_t; ++t, ++g) {
          const uint32_t s = g % PT_STAGES;
          const uint32_t ph = (g / PT_STAGES) & 1u;
          pt_wait(&b_empty[s], ph ^ 1u);
          mbar_arrive_expect_tx(&b_full[s], PT_B_BYTES);
          uint8_t* dst = sB + (size_t)s * PT_B_BYTES;
          tma_bulk_g2s(dst, w.b_exp + (size_t)t * (PT_N * 128), PT_N * 128, &b_full[s]);
          tma_bulk_g2s(dst + PT_N * 128, w.b_exp + (size_t)w.b_rows_pad * 128 + (size_t)t * (PT_N * 128),
                       PT_N * 128, &b_full[s]);
        }
      }
    }
  } else if (warp == PT_EPI_WARPS + 1) {
    // ===== MMA issuer =====
    if (lane == 0) {
      uint32_t g = 0, item = 0;
      const uint32_t a_addr = smem_u32(sA);
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++item) {
        const PtWork w = work[pt_find_work(work, n_work, tile)];
        const int n_t = (w.p2 + PT_N - 1) / PT_N;
        pt_wait(a_full, item & 1u);
        for (int t = 0; t < n_t; ++t, ++g) {
          const uint32_t s = g % PT_STAGES;
          const uint32_t ph = (g / PT_STAGES) & 1u;
          const uint32_t buf = g % PT_ACC;
          pt_wait(&b_full[s], ph);
          pt_wait(&acc_empty[buf], ((g / PT_ACC) & 1u) ^ 1u);
          tc_fence_after();
          const uint32_t b_addr = smem_u32(sB + (size_t)s * PT_B_BYTES);
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            const uint64_t ad = tc_smem_desc(a_addr + (k >> 2) * (PT_M * 128) + (k & 3) * 32);
            const uint64_t bd = tc_smem_desc(b_addr + (k >> 2) * (PT_N * 128) + (k & 3) * 32);
            pt_mma_tf32(tmem_base + buf * PT_N, ad, bd, PT_IDESC, k > 0 ? 1u : 0u);
          }
          tc_commit(&b_empty[s]);
          tc_commit(&acc_full[buf]);
        }
        tc_commit(a_empty);
      }
    }
  } else {
    // ===== epilogue warpgroups: one row per thread, wg 0 the columns [0, 128) of every tile, wg 1 [128, 256) =====
    const uint32_t wg = (uint32_t)warp >> 2;
    const int row_in_tile = (warp & 3) * 32 + lane;
    const uint32_t taddr0 = tmem_base + wg * PT_HALF + ((uint32_t)((warp & 3) * 32) << 16);
    uint32_t g = 0, item = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++item) {
      const PtWork w = work[pt_find_work(work, n_work, tile)];
      const int n_t = (w.p2 + PT_N - 1) / PT_N;
      const int row = (tile - w.tile_begin) * PT_M + row_in_tile;
      const float f = fmaxf(*w.f_a, *w.f_b);
      if (!(f < PT_F_MAX) && tid == 0) atomicOr(flags, PT_FLAG_RANGE);
      // (the floor keeps the columns of an all-zero bank candidates of each other -- exact ties -- instead of none)
      const float e3 = fmaxf(3.0f * PT_ERR_COEF * f * f, 1e-30f);
      float theta = row < w.p1 ? INFINITY : -INFINITY;     // a padding row takes nothing
      volatile float* row_slots = share + row_in_tile;   // [wg * PT_M]
      volatile float* my_slot = row_slots + wg * PT_M;
      for (int t = 0; t < n_t; ++t, ++g) {
        const uint32_t buf = g % PT_ACC;
        __syncwarp();                        // the rare path of the previous tile diverges
#pragma unroll
        for (int o = 1; o < PT_WGS; ++o) theta = fminf(theta, row_slots[((wg + o) % PT_WGS) * PT_M]);
        pt_wait(&acc_full[buf], (g / PT_ACC) & 1u);
        tc_fence_after();
        const uint32_t taddr = taddr0 + buf * PT_N;
        uint32_t v[PT_HALF];
#if PT_EXPERIMENT != 1
        pt_ld32<0>(taddr, v);
        pt_ld32<32>(taddr, v);
        if constexpr (PT_HALF == 128) {
          pt_ld32<64>(taddr, v);
          pt_ld32<96>(taddr, v);
        }
        pt_wait_ld32<0>(v);
        pt_wait_ld32<32>(v);
        if constexpr (PT_HALF == 128) {
          pt_wait_ld32<64>(v);
          pt_wait_ld32<96>(v);
        }
#endif
        tc_release(&acc_empty[buf], lane);   // the accumulator goes back as soon as it sits in registers
        const int col0 = t * PT_N + (int)wg * PT_HALF;
#if PT_EXPERIMENT == 0
        pt_chunk<0>(v, theta, e3, w, row, col0, cand_n, seg, seg_cap, flags, my_slot);
        pt_chunk<32>(v, theta, e3, w, row, col0, cand_n, seg, seg_cap, flags, my_slot);
        if constexpr (PT_HALF == 128) {
          pt_chunk<64>(v, theta, e3, w, row, col0, cand_n, seg, seg_cap, flags, my_slot);
          pt_chunk<96>(v, theta, e3, w, row, col0, cand_n, seg, seg_cap, flags, my_slot);
        }
#elif PT_EXPERIMENT == 2
        if (__uint_as_float(v[0]) == 12345.678f && __uint_as_float(v[PT_HALF - 1]) == 1.0f) *my_slot = (float)col0;   // keeps the loads alive
#endif
      }
      // The row's FINAL theta (the smallest over its threads) for the exact part: a candidate taken early,
      // with D at or above it, is ruled out like a column that was never taken and needs no FP64 evaluation.
      if (row < w.p1) atomicMin(&row_theta[w.out_row + row], pt_ordered(theta));
      // every thread resets its slot between two meetings of the warpgroups, so that none of them reads a
      // theta of the previous row there
      asm volatile("bar.sync 1, %0;" ::"n"(32 * PT_EPI_WARPS) : "memory");
      *my_slot = INFINITY;
      asm volatile("bar.sync 1, %0;" ::"n"(32 * PT_EPI_WARPS) : "memory");
    }
    if (tid == 0) cand_counts[blockIdx.x] = min(*cand_n, seg_cap);
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 9) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem_base) : "memory");
  }
}

// ---- exact part ------------------------------------------------------------------------------------
constexpr unsigned long long PT_KEY_NONE = 0x7FF0000000000000ull;   // +inf

__global__ void __launch_bounds__(256)
poly_cand_init_kernel(unsigned long long* __restrict__ row_key, int* __restrict__ argmin,
                      unsigned* __restrict__ row_theta, long long rows) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < rows) {
    row_key[i] = PT_KEY_NONE;
    argmin[i] = 0x7FFFFFFF;
    row_theta[i] = 0xFFFFFFFFu;
  }
}

// Which candidates need the FP64 evaluation: those whose D lies below the row's final theta (a candidate whose D
// the screen did not keep -- NaN -- always does).  Their pool positions go to a dense list (one atomicAdd per
// warp), so that the evaluation below runs with full warps: at C4 nine of ten candidates are ruled out, and a
// warp that evaluates one lane in eight costs as much as a full one.
__global__ void __launch_bounds__(256)
poly_cand_filter_kernel(const PolyPair* __restrict__ pairs, const PtCand* __restrict__ pool, unsigned seg_cap,
                        const unsigned* __restrict__ cand_counts, const unsigned* __restrict__ row_theta,
                        unsigned* __restrict__ keep, unsigned* __restrict__ n_keep) {
  constexpr unsigned FULL = 0xffffffffu;
  const unsigned seg = blockIdx.y;
  const unsigned n = cand_counts[seg];
  const int lane = threadIdx.x & 31;
  const unsigned stride = gridDim.x * blockDim.x;
  for (unsigned i0 = blockIdx.x * blockDim.x + (threadIdx.x & ~31u); i0 < n; i0 += stride) {   // warp-uniform
    const unsigned i = i0 + lane;
    bool kept = false;
    unsigned at = 0;
    if (i < n) {
      at = seg * seg_cap + i;
      const PtCand c = pool[at];
      kept = !(c.d >= pt_unordered(row_theta[pairs[c.pair].out_row + c.row]));
    }
    const unsigned m = __ballot_sync(FULL, kept);
    if (m == 0u) continue;
    unsigned base = 0u;
    if (lane == 0) base = atomicAdd(n_keep, (unsigned)__popc(m));
    base = __shfl_sync(FULL, base, 0);
    if (kept) keep[base + __popc(m & ((1u << lane) - 1u))] = at;
  }
}

// one thread per KEPT candidate; stats[0] = max abs(|a|^2 + D - v) / F^2 as float bits
__global__ void __launch_bounds__(256)
poly_cand_eval_kernel(const PolyPair* __restrict__ pairs, const PtWork* __restrict__ work,
                      const PtCand* __restrict__ pool, const unsigned* __restrict__ keep,
                      const unsigned* __restrict__ n_keep, const double* const* __restrict__ na2_of_pair,
                      double* __restrict__ keep_sv, unsigned long long* __restrict__ row_key,
                      unsigned* __restrict__ stats, int* __restrict__ flags,
                      unsigned long long* __restrict__ executed_flops) {
  const unsigned n = *n_keep;
  unsigned evaluated = 0;
  float worst = 0.0f;
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const PtCand c = pool[keep[i]];
    const PolyPair pp = pairs[c.pair];
    double a[15];
#pragma unroll
    for (int k = 0; k < 15; ++k) a[k] = pp.f1[(long long)c.row * 15 + k];
    const double v = poly_dist2_f64<0>(a, pp.f2 + (long long)c.col * 15);
    ++evaluated;
    const double s_screen = na2_of_pair[c.pair][c.row] + (double)c.d;
    const PtWork w = work[c.pair];
    const float f = fmaxf(*w.f_a, *w.f_b);
    const double e = (double)PT_ERR_COEF * (double)f * (double)f;
    const double diff = fabs(s_screen - v);
    if (diff > e) atomicOr(flags, PT_FLAG_ERROR);            // NaN differences (NaN features, D not kept) are not errors
    if (f > 0.0f) worst = fmaxf(worst, __double2float_ru(diff / ((double)f * (double)f)));
    const double sv = sqrt(v);
    const bool valid = v < INFINITY;                           // what `v < thr` admits from thr = inf on
    keep_sv[i] = valid ? sv : INFINITY;
    if (valid) atomicMin(&row_key[pp.out_row + c.row], (unsigned long long)__double_as_longlong(sv));
  }
  worst = __uint_as_float(__reduce_max_sync(0xffffffffu, __float_as_uint(worst)));
  evaluated = __reduce_add_sync(0xffffffffu, evaluated);
  if ((threadIdx.x & 31) == 0) {
    if (worst > 0.0f) atomicMax(&stats[0], __float_as_uint(worst));
    if (evaluated) atomicAdd(executed_flops, 45ull * evaluated);
  }
}

__global__ void __launch_bounds__(256)
poly_cand_argmin_kernel(const PolyPair* __restrict__ pairs, const PtCand* __restrict__ pool,
                        const unsigned* __restrict__ keep, const unsigned* __restrict__ n_keep,
                        const double* __restrict__ keep_sv, const unsigned long long* __restrict__ row_key,
                        int* __restrict__ argmin) {
  const unsigned n = *n_keep;
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const double sv = keep_sv[i];
    if (!(sv < INFINITY)) continue;
    const PtCand c = pool[keep[i]];
    const long long r = pairs[c.pair].out_row + c.row;
    if ((unsigned long long)__double_as_longlong(sv) == row_key[r]) atomicMin(&argmin[r], c.col);
  }
}

__global__ void __launch_bounds__(256)
poly_cand_finish_kernel(const unsigned long long* __restrict__ row_key, int* __restrict__ argmin,
                        double* __restrict__ mindist, long long rows) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < rows) {
    const unsigned long long k = row_key[i];
    mindist[i] = __longlong_as_double((long long)k);
    if (k == PT_KEY_NONE) argmin[i] = 0;
  }
}

}  // namespace ssp
