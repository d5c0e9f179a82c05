// Kernel 1 — all-pairs Hamming nearest neighbours on 256-bit ORB descriptors (sm_100a).
//
// Replaces cv.BFMatcher(NORM_HAMMING).knnMatch(k=2) + ratio test
// (custom_bruteforce_matcher.py:160-172) and cv.BFMatcher(NORM_HAMMING, crossCheck=True).match
// (custom_bruteforce_matcher.py:193-199) for a whole list of image pairs.
//
// POPC variant.  One CTA owns a tile of HB_THREADS*R query rows of one directed work item
// (query image -> train image); every thread keeps R query descriptors (8 x b32 each) in
// registers.  A dedicated producer warp streams the train image through a ring of shared-memory
// stages with TMA 1-D bulk copies (cp.async.bulk + mbarrier complete_tx); the 8 consumer warps
// read each train row as two broadcast LDS.128 and evaluate it against their R queries with
// LOP3 (xor / carry-save adders) + POPC.  The running best-2 per query lives in registers as a
// packed key (distance << 22 | train index): an unsigned min on the key IS the
// (distance, lowest index) order OpenCV's stable insertion produces, so no tie handling is
// needed anywhere.  Finalize kernels apply the ratio test / mutual-nearest test on the device
// and compact the survivors in ascending query order.
#include <cstdlib>

#include "common.cuh"

namespace ssp {

constexpr int HB_THREADS = 256;          // consumer threads per CTA
constexpr int HB_CHUNK = 256;            // train rows per shared-memory stage (8 KB)
constexpr int HB_STAGES = 4;
constexpr uint32_t KEY_SHIFT = 22;       // distance <= 256 -> key < 2^31
constexpr uint32_t KEY_IDX_MASK = (1u << KEY_SHIFT) - 1u;
constexpr uint32_t KEY_NONE = 0xFFFFFFFFu;

__device__ __forceinline__ void csa(uint32_t a, uint32_t b, uint32_t c, uint32_t& sum,
                                    uint32_t& carry) {
  // one LOP3 each: 0x96 = a^b^c, 0xE8 = majority(a,b,c)
  asm("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(sum) : "r"(a), "r"(b), "r"(c));
  asm("lop3.b32 %0, %1, %2, %3, 0xE8;" : "=r"(carry) : "r"(a), "r"(b), "r"(c));
}

// Hamming distance of two 256-bit rows.  NPOPC selects how many POPC instructions are spent:
//   8: xor + popc per word (the textbook form, 8 LOP3 + 8 POPC + adds)
//   4: Harley-Seal carry-save compression, 8 LOP3 (xor) + 8 LOP3 (4 CSA) + 4 POPC
template <int NPOPC>
__device__ __forceinline__ uint32_t ham256(const uint32_t (&q)[8], const uint4 t0, const uint4 t1) {
  const uint32_t x0 = q[0] ^ t0.x, x1 = q[1] ^ t0.y, x2 = q[2] ^ t0.z, x3 = q[3] ^ t0.w;
  const uint32_t x4 = q[4] ^ t1.x, x5 = q[5] ^ t1.y, x6 = q[6] ^ t1.z, x7 = q[7] ^ t1.w;
  if (NPOPC == 8) {
    return (__popc(x0) + __popc(x1) + __popc(x2)) + (__popc(x3) + __popc(x4) + __popc(x5)) +
           (__popc(x6) + __popc(x7));
  } else {
    uint32_t s0, c0, s1, c1, s2, c2, s3, c3;
    csa(x0, x1, x2, s0, c0);
    csa(x3, x4, x5, s1, c1);
    csa(s0, s1, x6, s2, c2);   // ones: s2, x7
    csa(c0, c1, c2, s3, c3);   // twos: s3 ; fours: c3
    return (__popc(s2) + __popc(x7)) + 2u * __popc(s3) + 4u * __popc(c3);
  }
}

template <int R, int NPOPC, int K>
__global__ void __launch_bounds__(HB_THREADS + 32)
hamming_popc_kernel(const HammingWork* __restrict__ work, int4* __restrict__ best2) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  uint4* stage_buf = reinterpret_cast<uint4*>(smem_raw);
  uint64_t* full = reinterpret_cast<uint64_t*>(smem_raw + (size_t)HB_STAGES * HB_CHUNK * 32);
  uint64_t* empty = full + HB_STAGES;

  const HammingWork w = work[blockIdx.x];
  const int tile_base = blockIdx.y * (HB_THREADS * R);
  if (tile_base >= w.nq) return;  // uniform per CTA

  const int tid = threadIdx.x;
  const int warp = tid >> 5;
  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < HB_STAGES; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], HB_THREADS / 32);
    }
    fence_mbar_init();
  }
  __syncthreads();

  const int n_chunks = (w.nt + HB_CHUNK - 1) / HB_CHUNK;

  if (warp == HB_THREADS / 32) {
    // ---- producer warp: one elected lane feeds the ring -----------------------------------
    if ((tid & 31) == 0) {
      for (int c = 0; c < n_chunks; ++c) {
        const int s = c % HB_STAGES;
        const uint32_t ph = (uint32_t)(c / HB_STAGES) & 1u;
        if (c >= HB_STAGES) mbar_wait(&empty[s], ph ^ 1u);
        const int rows = min(HB_CHUNK, w.nt - c * HB_CHUNK);
        const uint32_t bytes = (uint32_t)rows * 32u;
        mbar_arrive_expect_tx(&full[s], bytes);
        tma_bulk_g2s(stage_buf + (size_t)s * HB_CHUNK * 2, w.t + (size_t)c * HB_CHUNK * 2, bytes,
                     &full[s]);
      }
    }
    return;
  }

  // ---- consumers --------------------------------------------------------------------------
  uint32_t q[R][8];
  uint32_t b1[R], b2[R];
#pragma unroll
  for (int r = 0; r < R; ++r) {
    const int row = tile_base + r * HB_THREADS + tid;
    uint4 a = make_uint4(0, 0, 0, 0), b = a;
    if (row < w.nq) {
      a = __ldg(w.q + (size_t)row * 2);
      b = __ldg(w.q + (size_t)row * 2 + 1);
    }
    q[r][0] = a.x; q[r][1] = a.y; q[r][2] = a.z; q[r][3] = a.w;
    q[r][4] = b.x; q[r][5] = b.y; q[r][6] = b.z; q[r][7] = b.w;
    b1[r] = KEY_NONE;
    b2[r] = KEY_NONE;
  }

  for (int c = 0; c < n_chunks; ++c) {
    const int s = c % HB_STAGES;
    const uint32_t ph = (uint32_t)(c / HB_STAGES) & 1u;
    mbar_wait(&full[s], ph);
    const uint4* tb = stage_buf + (size_t)s * HB_CHUNK * 2;
    const int rows = min(HB_CHUNK, w.nt - c * HB_CHUNK);
    const uint32_t base = (uint32_t)c * HB_CHUNK;
#pragma unroll 2
    for (int j = 0; j < rows; ++j) {
      const uint4 t0 = tb[2 * j];
      const uint4 t1 = tb[2 * j + 1];
      const uint32_t idx = base + (uint32_t)j;
#pragma unroll
      for (int r = 0; r < R; ++r) {
        const uint32_t d = ham256<NPOPC>(q[r], t0, t1);
        const uint32_t key = d * (1u << KEY_SHIFT) + idx;
        if (K == 2) b2[r] = min(b2[r], max(b1[r], key));
        b1[r] = min(b1[r], key);
      }
    }
    __syncwarp();
    if ((tid & 31) == 0) mbar_arrive(&empty[s]);
  }

#pragma unroll
  for (int r = 0; r < R; ++r) {
    const int row = tile_base + r * HB_THREADS + tid;
    if (row < w.nq) {
      int4 o;
      o.x = (b1[r] == KEY_NONE) ? -1 : (int)(b1[r] & KEY_IDX_MASK);
      o.y = (b1[r] == KEY_NONE) ? -1 : (int)(b1[r] >> KEY_SHIFT);
      o.z = (K == 1 || b2[r] == KEY_NONE) ? -1 : (int)(b2[r] & KEY_IDX_MASK);
      o.w = (K == 1 || b2[r] == KEY_NONE) ? -1 : (int)(b2[r] >> KEY_SHIFT);
      best2[w.out_row + row] = o;
    }
  }
}

// ---- finalize: survivor test, per-pair count, scan over pairs, ordered compaction ----------

// internal finalize mode: cross-check whose reverse direction sits in the column-key array
constexpr int SSP_MODE_CROSS_FUSED = 2;
constexpr uint32_t COL_KEY_ROW_MASK = (1u << 18) - 1u;

__device__ __forceinline__ bool survives(const HammingPair& p, const int4* __restrict__ best2,
                                         const uint32_t* __restrict__ col_key, int q, int mode, double ratio,
                                         int4& rec) {
  rec = best2[p.fwd_row + q];
  if (mode == SSP_MODE_KNN2_RATIO) {
    // `m.distance < ratio * n.distance` exactly as Python evaluates it (custom_bruteforce_matcher.py:
    // 164): both distances are small integers held in floats, the product and the comparison are
    // float64 -- one correctly rounded multiplication, no FMA involved
    return rec.z >= 0 && (double)rec.y < __dmul_rn(ratio, (double)rec.w);
  }
  if (rec.x < 0) return false;
  if (mode == SSP_MODE_CROSS_FUSED) return (col_key[p.rev_row + rec.x] & COL_KEY_ROW_MASK) == (uint32_t)q;
  return best2[p.rev_row + rec.x].x == q;
}

// fused cross-check: thresholds of real train rows start at "anything gets through", those of the
// pad rows of the last tile at "nothing does" (hamming_tensor.cu, tc_cols)
__global__ void __launch_bounds__(256)
col_thr_init_kernel(const HammingPair* __restrict__ pairs, int16_t* __restrict__ thr) {
  const HammingPair p = pairs[blockIdx.x];
  const int nt_pad = (p.nt + 255) / 256 * 256;
  for (int t = blockIdx.y * blockDim.x + threadIdx.x; t < nt_pad; t += gridDim.y * blockDim.x)
    thr[p.rev_row + t] = (t < p.nt) ? (int16_t)16382 : (int16_t)-16384;
}

// ---- cross-check, tensor variant: the reverse pass runs over the CLAIMED train rows only --------
// A couple (q, t) survives the cross-check iff t is q's nearest neighbour and q is t's, so the
// reverse direction is needed for exactly those rows of the train image that are somebody's
// nearest neighbour (~63 % of them on unrelated descriptors).  cc_mark flags them from the forward
// records, cc_slots lays them out as the sparse query tiles of HammingWork::slot_src.
struct CrossAux {
  long long fwd_row;    // forward best2 rows of the pair
  long long mark_off;   // first mark byte of the pair (nt of them)
  long long slot_off;   // first slot of the pair
  int nq, nt;
};

// mark[t] = smallest distance at which some query row claims train row t (CC_UNCLAIMED if none):
// the reverse search of t cannot end above it, so it starts with that threshold already settled.
constexpr int CC_UNCLAIMED = 0x7F7F7F7F;   // what cudaMemset(0x7F) leaves
__global__ void __launch_bounds__(256)
cc_mark_kernel(const CrossAux* __restrict__ aux, const int4* __restrict__ best2, int* __restrict__ mark) {
  const CrossAux a = aux[blockIdx.x];
  for (int q = blockIdx.y * blockDim.x + threadIdx.x; q < a.nq; q += gridDim.y * blockDim.x) {
    const int4 rec = best2[a.fwd_row + q];
    if (rec.x >= 0) atomicMin(&mark[a.mark_off + rec.x], rec.y);
  }
}

// block (pair, c): the marked rows t = c (mod 8) in ascending order go to slots c, c + 8, c + 16, ...
__global__ void __launch_bounds__(256)
cc_slots_kernel(const CrossAux* __restrict__ aux, const int* __restrict__ mark, int* __restrict__ slots,
                int* __restrict__ cls_count) {
  const CrossAux a = aux[blockIdx.x];
  const int c = blockIdx.y;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  __shared__ int warp_cnt[8];
  __shared__ int run;
  if (threadIdx.x == 0) run = 0;
  __syncthreads();
  const int n_c = (a.nt - c + 7) / 8;   // rows of this class
  for (int base = 0; base < n_c; base += 256) {
    const int i = base + threadIdx.x;
    const int t = c + 8 * i;
    const int claimed = (i < n_c) ? mark[a.mark_off + t] : CC_UNCLAIMED;
    const bool keep = claimed != CC_UNCLAIMED;
    const unsigned ballot = __ballot_sync(0xffffffffu, keep);
    if (lane == 0) warp_cnt[wid] = __popc(ballot);
    __syncthreads();
    int before = 0, total = 0;
#pragma unroll
    for (int w = 0; w < 8; ++w) {
      const int n = warp_cnt[w];
      before += (w < wid) ? n : 0;
      total += n;
    }
    // slot value: row | claimed distance << SSP_SLOT_DIST_SHIFT (rows < 2^18, distances <= 256)
    if (keep)
      slots[a.slot_off + (long long)(run + before + __popc(ballot & ((1u << lane) - 1u))) * 8 + c] =
          t | (claimed << SSP_SLOT_DIST_SHIFT);
    __syncthreads();
    if (threadIdx.x == 0) run += total;
    __syncthreads();
  }
  if (threadIdx.x == 0) cls_count[blockIdx.x * 8 + c] = run;
}

constexpr int FIN_THREADS = 512;

__global__ void __launch_bounds__(FIN_THREADS)
hamming_count_kernel(const HammingPair* __restrict__ pairs, const int4* __restrict__ best2,
                     const uint32_t* __restrict__ col_key, int mode, double ratio, int* __restrict__ counts) {
  const HammingPair p = pairs[blockIdx.x];
  int local = 0;
  for (int q = threadIdx.x; q < p.nq; q += FIN_THREADS) {
    int4 rec;
    local += survives(p, best2, col_key, q, mode, ratio, rec) ? 1 : 0;
  }
  __shared__ int warp_sum[FIN_THREADS / 32];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) local += __shfl_xor_sync(0xffffffffu, local, o);
  if ((threadIdx.x & 31) == 0) warp_sum[threadIdx.x >> 5] = local;
  __syncthreads();
  if (threadIdx.x == 0) {
    int tot = 0;
    for (int i = 0; i < FIN_THREADS / 32; ++i) tot += warp_sum[i];
    counts[blockIdx.x] = tot;
  }
}

__global__ void __launch_bounds__(FIN_THREADS)
hamming_compact_kernel(const HammingPair* __restrict__ pairs, const int4* __restrict__ best2,
                       const uint32_t* __restrict__ col_key, int mode, double ratio,
                       const long long* __restrict__ offsets,
                       int* __restrict__ mq, int* __restrict__ mt, int* __restrict__ md) {
  const HammingPair p = pairs[blockIdx.x];
  __shared__ int warp_cnt[FIN_THREADS / 32];
  __shared__ long long run;
  if (threadIdx.x == 0) run = offsets[blockIdx.x];
  __syncthreads();
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  for (int base = 0; base < p.nq; base += FIN_THREADS) {
    const int q = base + threadIdx.x;
    int4 rec = make_int4(0, 0, 0, 0);
    const bool keep = (q < p.nq) && survives(p, best2, col_key, q, mode, ratio, rec);
    const unsigned ballot = __ballot_sync(0xffffffffu, keep);
    if (lane == 0) warp_cnt[wid] = __popc(ballot);
    __syncthreads();
    int before = 0, total = 0;
#pragma unroll
    for (int i = 0; i < FIN_THREADS / 32; ++i) {
      const int c = warp_cnt[i];
      before += (i < wid) ? c : 0;
      total += c;
    }
    if (keep) {
      const long long pos = run + before + __popc(ballot & ((1u << lane) - 1u));
      mq[pos] = q;
      mt[pos] = rec.x;
      md[pos] = rec.y;
    }
    __syncthreads();
    if (threadIdx.x == 0) run += total;
    __syncthreads();
  }
}

// ---- host side ------------------------------------------------------------------------------

template <int R, int NPOPC, int K>
static cudaError_t launch_popc(ssp_ctx* ctx, int n_work, int max_nq) {
  const size_t smem = (size_t)HB_STAGES * HB_CHUNK * 32 + 2 * HB_STAGES * sizeof(uint64_t);
  auto kern = hamming_popc_kernel<R, NPOPC, K>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  const int tiles = (max_nq + HB_THREADS * R - 1) / (HB_THREADS * R);
  dim3 grid((unsigned)n_work, (unsigned)tiles);
  kern<<<grid, HB_THREADS + 32, smem, ctx->stream>>>(ctx->h_work.as<HammingWork>(),
                                                       ctx->h_best2.as<int4>());
  return cudaGetLastError();
}

// implemented in hamming_tensor.cu
cudaError_t launch_hamming_tensor(ssp_ctx* ctx, const std::vector<HammingWork>& work, int k,
                                  std::string* why_not, bool wait);

}  // namespace ssp

namespace ssp {

// How the tensor variant does the cross-check (environment variable SSP_B200_CROSSCHECK):
//   sparse (default) forward launch, then a reverse launch over the claimed rows only (1.64 passes)
//   fused            one launch: the column direction comes out of the forward accumulators (exact, but
//                    ~1.5 column events per warp and tile make it slower than two passes: DESIGN.md)
//   dense            both directions as dense work items of one launch (what the POPC variants do)
enum CrossForm { CC_FUSED, CC_SPARSE, CC_DENSE };
static CrossForm cross_form() {   // read per call: tests and A/B runs switch it between calls
  const char* d = getenv("SSP_B200_CROSSCHECK_DENSE");
  if (d && d[0] == '1') return CC_DENSE;
  const char* e = getenv("SSP_B200_CROSSCHECK");
  if (e && e[0] == 'f') return CC_FUSED;
  if (e && e[0] == 'd') return CC_DENSE;
  return CC_SPARSE;
}

static int hamming_match_impl(ssp_ctx* ctx, const int32_t* pairs, int n_pairs, int mode, double ratio,
                              int variant, int64_t* match_offsets) {
  if (!ctx) return SSP_ERR_INVALID;
  if (int rc = use_device(ctx)) return rc;
  SSP_REQUIRE(ctx, ctx->d_desc != nullptr || (!ctx->row_off.empty() && ctx->row_off.back() == 0),
              SSP_ERR_STATE, "ssp_hamming_match_pairs: no descriptor bank uploaded");
  SSP_REQUIRE(ctx, !ctx->row_off.empty(), SSP_ERR_STATE,
              "ssp_hamming_match_pairs: no descriptor bank uploaded");
  SSP_REQUIRE(ctx, n_pairs >= 0 && (n_pairs == 0 || pairs != nullptr), SSP_ERR_INVALID,
              "ssp_hamming_match_pairs: bad pair list");
  SSP_REQUIRE(ctx, match_offsets != nullptr, SSP_ERR_INVALID,
              "ssp_hamming_match_pairs: match_offsets is NULL");
  SSP_REQUIRE(ctx, mode == SSP_MODE_KNN2_RATIO || mode == SSP_MODE_CROSSCHECK, SSP_ERR_INVALID,
              "ssp_hamming_match_pairs: unknown mode %d", mode);
  SSP_REQUIRE(ctx, variant >= SSP_HAMMING_AUTO && variant <= SSP_HAMMING_TENSOR, SSP_ERR_INVALID,
              "ssp_hamming_match_pairs: unknown variant %d", variant);
  if (mode == SSP_MODE_KNN2_RATIO)
    SSP_REQUIRE(ctx, ratio > 0.0 && ratio == ratio, SSP_ERR_INVALID,
                "ssp_hamming_match_pairs: ratio must be positive");
  const int n_images = (int)ctx->row_off.size() - 1;
  NvtxRange range_all("ssp_hamming_match_pairs");

  ctx->h_n_matches = -1;
  ctx->h_launches = 0;
  ctx->h_main_ms = 0.f;
  ctx->h_comparisons = 0;
  ctx->h_match_offsets.assign((size_t)n_pairs + 1, 0);
  if (n_pairs == 0) {
    match_offsets[0] = 0;
    ctx->h_n_matches = 0;
    ctx->h_total_rows = 0;
    return SSP_OK;
  }

  // directed work items + per-pair finalize records
  const bool cross = (mode == SSP_MODE_CROSSCHECK);
  int v = variant;
  if (v == SSP_HAMMING_AUTO) v = SSP_HAMMING_TENSOR;   // fastest at every size measured
  const bool tensor = (v == SSP_HAMMING_TENSOR);
  const CrossForm form = (cross && tensor) ? cross_form() : CC_DENSE;
  const bool fused = cross && form == CC_FUSED;
  const bool sparse_rev = cross && form == CC_SPARSE;
  std::vector<HammingWork> work, rev_work;
  std::vector<HammingPair> hp((size_t)n_pairs);
  work.reserve((size_t)n_pairs * (cross ? 2 : 1));
  const uint4* bank = reinterpret_cast<const uint4*>(ctx->d_desc);
  long long fwd_rows = 0, rows = 0;
  int max_nq = 0, max_nt = 1;
  for (int p = 0; p < n_pairs; ++p) {
    const int qi = pairs[2 * p], ti = pairs[2 * p + 1];
    SSP_REQUIRE(ctx, qi >= 0 && qi < n_images && ti >= 0 && ti < n_images, SSP_ERR_INVALID,
                "ssp_hamming_match_pairs: pair %d = (%d, %d) out of range [0, %d)", p, qi, ti,
                n_images);
    // cv::BFMatcher packs (image, row) into one int and asserts rows < 2^18 per train set
    // (IMGIDX_SHIFT); the packed best-2 keys here leave 22 bits, the limit kept is OpenCV's.
    // Cross-check matches both directions, so either image is a train set.
    const long long nq_p = ctx->row_off[qi + 1] - ctx->row_off[qi];
    const long long nt_p = ctx->row_off[ti + 1] - ctx->row_off[ti];
    SSP_REQUIRE(ctx, nt_p < (1LL << 18) && (!cross || nq_p < (1LL << 18)), SSP_ERR_INVALID,
                "ssp_hamming_match_pairs: pair %d = (%d, %d): a train set of %lld rows exceeds "
                "cv::BFMatcher's limit of 2^18 - 1 rows per image", p, qi, ti,
                cross && nq_p > nt_p ? nq_p : nt_p);
    fwd_rows += nq_p;
  }
  long long rev_cursor = fwd_rows;   // two-pass forms: reverse best2 rows follow the forward ones
  long long col_cursor = 0;          // fused form: entries of the column-key array
  for (int p = 0; p < n_pairs; ++p) {
    const int qi = pairs[2 * p], ti = pairs[2 * p + 1];
    const int nq = (int)(ctx->row_off[qi + 1] - ctx->row_off[qi]);
    const int nt = (int)(ctx->row_off[ti + 1] - ctx->row_off[ti]);
    HammingWork f;
    f.q = bank + ctx->row_off[qi] * 2;
    f.t = bank + ctx->row_off[ti] * 2;
    f.nq = nq;
    f.nt = nt;
    f.out_row = rows;
    f.qi = qi;
    f.ti = ti;
    hp[p].fwd_row = rows;
    hp[p].rev_row = -1;
    hp[p].nq = nq;
    hp[p].nt = nt;
    rows += nq;
    max_nq = nq > max_nq ? nq : max_nq;
    max_nt = nt > max_nt ? nt : max_nt;
    ctx->h_comparisons += (long long)nq * nt;
    if (fused) {
      hp[p].rev_row = col_cursor;               // 256-entry aligned: 16-byte aligned threshold rows
      col_cursor += ((long long)nt + 255) / 256 * 256;
    }
    if (nq > 0) work.push_back(f);              // col pointers are filled in below (fused)
    if (cross && !fused) {
      HammingWork r;
      r.q = f.t;
      r.t = f.q;
      r.nq = nt;
      r.nt = nq;
      r.out_row = rev_cursor;
      r.qi = ti;
      r.ti = qi;
      hp[p].rev_row = rev_cursor;
      rev_cursor += nt;
      if (nt > 0) (sparse_rev ? rev_work : work).push_back(r);
      max_nq = nt > max_nq ? nt : max_nq;
    }
  }
  const long long total_rows = (cross && !fused) ? rev_cursor : rows;
  ctx->h_total_rows = rows;

  SSP_CUDA_TRY(ctx, ctx->h_best2.ensure((size_t)(total_rows + 1) * sizeof(int4)));
  SSP_CUDA_TRY(ctx, ctx->h_pairs.ensure(hp.size() * sizeof(HammingPair)));
  SSP_CUDA_TRY(ctx, ctx->h_counts.ensure((size_t)n_pairs * sizeof(int)));
  SSP_CUDA_TRY(ctx, ctx->h_offsets.ensure(((size_t)n_pairs + 2) * sizeof(long long)));
  // survivors: at most one per forward query row, so the output arrays are sized up front and the
  // whole call needs ONE host synchronisation (at its end)
  SSP_CUDA_TRY(ctx, ctx->h_mq.ensure((size_t)(rows + 1) * sizeof(int)));
  SSP_CUDA_TRY(ctx, ctx->h_mt.ensure((size_t)(rows + 1) * sizeof(int)));
  SSP_CUDA_TRY(ctx, ctx->h_md.ensure((size_t)(rows + 1) * sizeof(int)));
  SSP_CUDA_TRY(ctx, ctx->h_tc_err.ensure(sizeof(int)));
  SSP_CUDA_TRY(ctx, cudaMemsetAsync(ctx->h_tc_err.p, 0, sizeof(int), ctx->stream));
  SSP_CUDA_TRY(ctx, cudaMemcpyAsync(ctx->h_pairs.p, hp.data(), hp.size() * sizeof(HammingPair),
                                    cudaMemcpyHostToDevice, ctx->stream));
  // rows of empty train sets are never written by a kernel: make them read as "no neighbour"
  SSP_CUDA_TRY(ctx, cudaMemsetAsync(ctx->h_best2.p, 0xFF, (size_t)(total_rows + 1) * sizeof(int4),
                                    ctx->stream));
  if (fused) {
    SSP_CUDA_TRY(ctx, ctx->h_col_key.ensure((size_t)(col_cursor + 1) * sizeof(uint32_t)));
    SSP_CUDA_TRY(ctx, ctx->h_col_thr.ensure((size_t)(col_cursor + 8) * sizeof(int16_t)));
    SSP_CUDA_TRY(ctx, cudaMemsetAsync(ctx->h_col_key.p, 0xFF, (size_t)(col_cursor + 1) * sizeof(uint32_t),
                                      ctx->stream));
    int by = (max_nt + 255) / 256;
    by = by > 32 ? 32 : by;
    col_thr_init_kernel<<<dim3((unsigned)n_pairs, (unsigned)by), 256, 0, ctx->stream>>>(
        ctx->h_pairs.as<HammingPair>(), ctx->h_col_thr.as<int16_t>());
    SSP_CUDA_TRY(ctx, cudaGetLastError());
    ctx->h_launches += 1;
    size_t wi = 0;
    for (int p = 0; p < n_pairs; ++p) {
      if (hp[(size_t)p].nq <= 0) continue;
      work[wi].col_key = ctx->h_col_key.as<uint32_t>() + hp[(size_t)p].rev_row;
      work[wi].col_thr = ctx->h_col_thr.as<int16_t>() + hp[(size_t)p].rev_row;
      ++wi;
    }
  }

  const int k = cross ? 1 : 2;
  float rev_fwd_ms = 0.f;   // sparse cross-check: kernel time of the forward launch (ev0/ev1 then hold the reverse one)
  bool timed = false;
  if (!work.empty()) {
    SSP_CUDA_TRY(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
    timed = true;
    if (tensor) {
      NvtxRange range_k("kernel 1 (tcgen05)");
      std::string why;
      cudaError_t e = launch_hamming_tensor(ctx, work, k, &why, /*wait=*/sparse_rev && !rev_work.empty());
      if (!why.empty()) return fail(ctx, SSP_ERR_INVALID, "tensor variant unavailable: %s", why.c_str());
      SSP_CUDA_TRY(ctx, e);
      if (sparse_rev && !rev_work.empty()) {
        float fwd_ms = 0.f;
        SSP_CUDA_TRY(ctx, cudaEventElapsedTime(&fwd_ms, ctx->ev0, ctx->ev1));  // the launch above has completed
        // claimed train rows of every pair -> sparse query tiles of its reverse item
        std::vector<CrossAux> aux((size_t)n_pairs);
        long long marks = 0, slots = 0;
        int max_q = 1;
        for (int p = 0; p < n_pairs; ++p) {
          aux[(size_t)p].fwd_row = hp[(size_t)p].fwd_row;
          aux[(size_t)p].mark_off = marks;
          aux[(size_t)p].slot_off = slots;
          aux[(size_t)p].nq = hp[(size_t)p].nq;
          aux[(size_t)p].nt = hp[(size_t)p].nt;
          marks += hp[(size_t)p].nt;
          slots += ((long long)hp[(size_t)p].nt + 127) / 128 * 128 + 256;
          max_q = hp[(size_t)p].nq > max_q ? hp[(size_t)p].nq : max_q;
        }
        SSP_CUDA_TRY(ctx, ctx->h_cc_aux.ensure(aux.size() * sizeof(CrossAux)));
        SSP_CUDA_TRY(ctx, ctx->h_cc_mark.ensure(((size_t)marks + 1) * sizeof(int)));
        SSP_CUDA_TRY(ctx, ctx->h_cc_slots.ensure((size_t)slots * sizeof(int)));
        SSP_CUDA_TRY(ctx, ctx->h_cc_cls.ensure((size_t)n_pairs * 8 * sizeof(int)));
        SSP_CUDA_TRY(ctx, cudaMemcpyAsync(ctx->h_cc_aux.p, aux.data(), aux.size() * sizeof(CrossAux),
                                          cudaMemcpyHostToDevice, ctx->stream));
        SSP_CUDA_TRY(ctx, cudaMemsetAsync(ctx->h_cc_mark.p, 0x7F, ((size_t)marks + 1) * sizeof(int), ctx->stream));
        SSP_CUDA_TRY(ctx, cudaMemsetAsync(ctx->h_cc_slots.p, 0xFF, (size_t)slots * sizeof(int), ctx->stream));
        int by = (max_q + 255) / 256;
        by = by > 64 ? 64 : by;
        cc_mark_kernel<<<dim3((unsigned)n_pairs, (unsigned)by), 256, 0, ctx->stream>>>(
            ctx->h_cc_aux.as<CrossAux>(), ctx->h_best2.as<int4>(), ctx->h_cc_mark.as<int>());
        SSP_CUDA_TRY(ctx, cudaGetLastError());
        cc_slots_kernel<<<dim3((unsigned)n_pairs, 8), 256, 0, ctx->stream>>>(
            ctx->h_cc_aux.as<CrossAux>(), ctx->h_cc_mark.as<int>(), ctx->h_cc_slots.as<int>(),
            ctx->h_cc_cls.as<int>());
        SSP_CUDA_TRY(ctx, cudaGetLastError());
        ctx->h_launches += 2;
        std::vector<int> cls((size_t)n_pairs * 8);
        SSP_CUDA_TRY(ctx, cudaMemcpyAsync(cls.data(), ctx->h_cc_cls.p, cls.size() * sizeof(int),
                                          cudaMemcpyDeviceToHost, ctx->stream));
        SSP_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
        size_t ri = 0;
        for (int p = 0; p < n_pairs; ++p) {
          if (hp[(size_t)p].nt <= 0) continue;   // no reverse item was made for this pair
          HammingWork& r = rev_work[ri++];
          int most = 0;
          for (int c = 0; c < 8; ++c) most = cls[(size_t)p * 8 + c] > most ? cls[(size_t)p * 8 + c] : most;
          r.slot_src = ctx->h_cc_slots.as<int>() + aux[(size_t)p].slot_off;
          r.q_tiles = (most + 15) / 16;   // 16 slots of every class per 128-slot tile
        }
        e = launch_hamming_tensor(ctx, rev_work, k, &why, /*wait=*/false);
        if (!why.empty()) return fail(ctx, SSP_ERR_INVALID, "tensor variant unavailable: %s", why.c_str());
        SSP_CUDA_TRY(ctx, e);
        rev_fwd_ms = fwd_ms;
      }
    } else {
      NvtxRange range_k("kernel 1 (popc)");
      SSP_CUDA_TRY(ctx, ctx->h_work.ensure(work.size() * sizeof(HammingWork)));
      SSP_CUDA_TRY(ctx, cudaMemcpyAsync(ctx->h_work.p, work.data(),
                                        work.size() * sizeof(HammingWork), cudaMemcpyHostToDevice,
                                        ctx->stream));
      cudaError_t e;
      const int nw = (int)work.size();
      if (v == SSP_HAMMING_POPC)
        e = (k == 2) ? launch_popc<4, 8, 2>(ctx, nw, max_nq) : launch_popc<4, 8, 1>(ctx, nw, max_nq);
      else
        e = (k == 2) ? launch_popc<4, 4, 2>(ctx, nw, max_nq) : launch_popc<4, 4, 1>(ctx, nw, max_nq);
      SSP_CUDA_TRY(ctx, e);
      ctx->h_launches += 1;
      SSP_CUDA_TRY(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
    }
  }

  {
    NvtxRange range_f("finalize (count, scan, compact)");
    const int fmode = fused ? SSP_MODE_CROSS_FUSED : mode;
    const uint32_t* ck = fused ? ctx->h_col_key.as<uint32_t>() : nullptr;
    hamming_count_kernel<<<n_pairs, FIN_THREADS, 0, ctx->stream>>>(
        ctx->h_pairs.as<HammingPair>(), ctx->h_best2.as<int4>(), ck, fmode, ratio, ctx->h_counts.as<int>());
    SSP_CUDA_TRY(ctx, cudaGetLastError());
    scan_counts_kernel<<<1, 1024, 0, ctx->stream>>>(ctx->h_counts.as<int>(), n_pairs,
                                                    ctx->h_offsets.as<long long>());
    SSP_CUDA_TRY(ctx, cudaGetLastError());
    hamming_compact_kernel<<<n_pairs, FIN_THREADS, 0, ctx->stream>>>(
        ctx->h_pairs.as<HammingPair>(), ctx->h_best2.as<int4>(), ck, fmode, ratio,
        ctx->h_offsets.as<long long>(), ctx->h_mq.as<int>(), ctx->h_mt.as<int>(), ctx->h_md.as<int>());
    SSP_CUDA_TRY(ctx, cudaGetLastError());
    ctx->h_launches += 3;
  }
  // offsets and the tensor kernel's error flag come back together: the call's only synchronisation
  int tc_flag = 0;
  SSP_CUDA_TRY(ctx, cudaMemcpyAsync(ctx->h_match_offsets.data(), ctx->h_offsets.p,
                                    ((size_t)n_pairs + 1) * sizeof(long long),
                                    cudaMemcpyDeviceToHost, ctx->stream));
  SSP_CUDA_TRY(ctx, cudaMemcpyAsync(&tc_flag, ctx->h_tc_err.p, sizeof(int), cudaMemcpyDeviceToHost,
                                    ctx->stream));
  SSP_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  if (tc_flag != 0)
    return fail(ctx, SSP_ERR_INVALID,
                "tensor variant unavailable: dynamic shared memory window is not 1024-byte aligned");
  const long long total = ctx->h_match_offsets[(size_t)n_pairs];
  if (timed) SSP_CUDA_TRY(ctx, cudaEventElapsedTime(&ctx->h_main_ms, ctx->ev0, ctx->ev1));
  ctx->h_main_ms += rev_fwd_ms;
  for (int p = 0; p <= n_pairs; ++p) match_offsets[p] = ctx->h_match_offsets[(size_t)p];
  ctx->h_n_matches = total;
  return SSP_OK;
}

}  // namespace ssp

extern "C" {

int ssp_hamming_match_pairs(ssp_ctx* ctx, const int32_t* pairs, int n_pairs, int mode,
                            int ratio_num, int ratio_den, int variant, int64_t* match_offsets) {
  if (!ctx) return SSP_ERR_INVALID;
  if (mode == SSP_MODE_KNN2_RATIO)
    SSP_REQUIRE(ctx, ratio_num > 0 && ratio_den > 0, SSP_ERR_INVALID,
                "ssp_hamming_match_pairs: ratio must be positive");
  const double ratio = (ratio_den > 0) ? (double)ratio_num / (double)ratio_den : 0.75;
  return ssp::hamming_match_impl(ctx, pairs, n_pairs, mode, ratio, variant, match_offsets);
}

int ssp_hamming_match_pairs_f64(ssp_ctx* ctx, const int32_t* pairs, int n_pairs, int mode,
                                double ratio, int variant, int64_t* match_offsets) {
  return ssp::hamming_match_impl(ctx, pairs, n_pairs, mode, ratio, variant, match_offsets);
}

int ssp_hamming_fetch_matches(ssp_ctx* ctx, int32_t* query_idx, int32_t* train_idx,
                              int32_t* distance, int64_t capacity, int mem_kind) {
  using namespace ssp;
  if (!ctx) return SSP_ERR_INVALID;
  if (int rc = use_device(ctx)) return rc;
  SSP_REQUIRE(ctx, ctx->h_n_matches >= 0, SSP_ERR_STATE,
              "ssp_hamming_fetch_matches: no completed ssp_hamming_match_pairs call");
  const long long n = ctx->h_n_matches;
  SSP_REQUIRE(ctx, capacity >= n, SSP_ERR_CAPACITY,
              "ssp_hamming_fetch_matches: capacity %lld < %lld matches", (long long)capacity, n);
  if (n == 0) return SSP_OK;
  const cudaMemcpyKind kind = from_device_kind(mem_kind);
  if (query_idx)
    SSP_CUDA_TRY(ctx, cudaMemcpyAsync(query_idx, ctx->h_mq.p, (size_t)n * 4, kind, ctx->stream));
  if (train_idx)
    SSP_CUDA_TRY(ctx, cudaMemcpyAsync(train_idx, ctx->h_mt.p, (size_t)n * 4, kind, ctx->stream));
  if (distance)
    SSP_CUDA_TRY(ctx, cudaMemcpyAsync(distance, ctx->h_md.p, (size_t)n * 4, kind, ctx->stream));
  SSP_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  return SSP_OK;
}

int ssp_hamming_fetch_best2(ssp_ctx* ctx, int32_t* best2, int64_t capacity_rows, int mem_kind) {
  using namespace ssp;
  if (!ctx) return SSP_ERR_INVALID;
  if (int rc = use_device(ctx)) return rc;
  SSP_REQUIRE(ctx, ctx->h_n_matches >= 0, SSP_ERR_STATE,
              "ssp_hamming_fetch_best2: no completed ssp_hamming_match_pairs call");
  SSP_REQUIRE(ctx, best2 != nullptr, SSP_ERR_INVALID, "ssp_hamming_fetch_best2: NULL buffer");
  SSP_REQUIRE(ctx, capacity_rows >= ctx->h_total_rows, SSP_ERR_CAPACITY,
              "ssp_hamming_fetch_best2: capacity %lld < %lld rows", (long long)capacity_rows,
              ctx->h_total_rows);
  if (ctx->h_total_rows == 0) return SSP_OK;
  SSP_CUDA_TRY(ctx, cudaMemcpyAsync(best2, ctx->h_best2.p, (size_t)ctx->h_total_rows * 16,
                                    from_device_kind(mem_kind), ctx->stream));
  SSP_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  return SSP_OK;
}

int ssp_hamming_last_stats(ssp_ctx* ctx, int* n_launches, float* main_kernel_ms,
                           int64_t* comparisons) {
  if (!ctx) return SSP_ERR_INVALID;
  if (n_launches) *n_launches = ctx->h_launches;
  if (main_kernel_ms) *main_kernel_ms = ctx->h_main_ms;
  if (comparisons) *comparisons = ctx->h_comparisons;
  return SSP_OK;
}

}  // extern "C"
