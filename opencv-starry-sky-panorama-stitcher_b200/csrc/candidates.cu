// Row N2 — candidate generation for the star-polygon path on the device (sm_100a).
//
// Replaces, behind the same host call, the parts of
// reduce_amount_of_keypoints_and_create_polygon_spanning_tuples_of_keypoints
// (custom_bf_polygon_matcher.py:261-443) and reduce_amount_of_keypoints_by_clustering (:137-259)
// that are plain data-parallel work:
//   * star_regions_kernel      the 12 overlapping border regions of cbpm:328-375 (strict < / >
//                              comparisons in float64, exactly the reference's expressions);
//   * cluster_brightest_kernel per cluster of a region keep the brightest star, the LAST of
//                              equally bright ones (cbpm:230-245: ascending stable sort, reversed);
//   * cand_kernel              all C(n_r, 5) combinations of every region's member list: the number of
//                              distinct TUPLES (what the reference's loop counts, cbpm:383, 407-413:
//                              itertools.combinations + set()) and the table of distinct star SETS
//                              (what gets measured), in a deterministic order.
// The Ward clustering itself (sklearn.cluster.AgglomerativeClustering, cbpm:175-181) stays on the
// host: its labels are an input here.  The candidate SET equals the reference's; the ORDER is
// region-major / lexicographic instead of CPython's set iteration order, which is why this path is
// opt-in on the host side (polygon order decides DMatch order).
#include "common.cuh"

namespace ssp {

constexpr int CG_REGIONS = 12;
constexpr int CG_MAX_MEMBERS = 32;   // members per region the enumeration accepts (reference: <= 15)
constexpr int CG_THREADS = 256;
constexpr int CG_MAX_STARS = 2048;   // star ids the duplicate test can index (reference: the 1000 brightest)

// cbpm:319-375.  Bit r of mask[i] = star i lies in region r, regions in the reference's dict order:
// left_top, left_bot, left_mid, right_top, right_bot, right_mid, top_left, top_right, top_mid,
// bot_left, bot_right, bot_mid.
__global__ void __launch_bounds__(256)
star_regions_kernel(const double* __restrict__ xy, int n, double img_w, double img_h,
                    unsigned short* __restrict__ mask) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double x = xy[2 * i], y = xy[2 * i + 1];
  const double vh = 0.70, vw = 0.33, hh = 0.33, hw = 0.50;   // cbpm:321-324
  // every product below is the reference's Python expression, evaluated in the same order
  const bool left = x < __dmul_rn(vw, img_w);
  const bool right = x > __dmul_rn(1.0 - vw, img_w);
  const bool v_top = y < __dmul_rn(vh, img_h);
  const bool v_bot = y > __dmul_rn(1.0 - vh, img_h);
  const bool v_mid = fabs(__dsub_rn(y, __dmul_rn(0.5, img_h))) < __ddiv_rn(__dmul_rn(vh, img_h), 2.0);
  const bool top = y < __dmul_rn(hh, img_h);
  const bool bot = y > __dmul_rn(1.0 - hh, img_h);
  const bool h_left = x < __dmul_rn(hw, img_w);
  const bool h_right = x > __dmul_rn(1.0 - hw, img_w);
  const bool h_mid = fabs(__dsub_rn(x, __dmul_rn(0.5, img_w))) < __ddiv_rn(__dmul_rn(hw, img_w), 2.0);
  unsigned m = 0;
  m |= (left && v_top) ? 1u << 0 : 0u;
  m |= (left && v_bot) ? 1u << 1 : 0u;
  m |= (left && v_mid) ? 1u << 2 : 0u;
  m |= (right && v_top) ? 1u << 3 : 0u;
  m |= (right && v_bot) ? 1u << 4 : 0u;
  m |= (right && v_mid) ? 1u << 5 : 0u;
  m |= (top && h_left) ? 1u << 6 : 0u;
  m |= (top && h_right) ? 1u << 7 : 0u;
  m |= (top && h_mid) ? 1u << 8 : 0u;
  m |= (bot && h_left) ? 1u << 9 : 0u;
  m |= (bot && h_right) ? 1u << 10 : 0u;
  m |= (bot && h_mid) ? 1u << 11 : 0u;
  mask[i] = (unsigned short)m;
}

// One block per (region, cluster label): among the members of the region that carry this label the
// brightest wins, the one listed LAST among equally bright ones.  members / labels are the region's
// stars in the order the host clustered them; survivor[region * max_labels + label] = star id or -1.
__global__ void __launch_bounds__(128)
cluster_brightest_kernel(const int* __restrict__ members, const int* __restrict__ labels,
                         const int* __restrict__ region_off, const double* __restrict__ bright,
                         int max_labels, int* __restrict__ survivor) {
  const int region = blockIdx.x, label = blockIdx.y;
  const int lo = region_off[region], hi = region_off[region + 1];
  // key = (brightness, position): maximum = brightest, last position among equals
  double best_b = -1.0;
  int best_pos = -1;
  for (int k = lo + threadIdx.x; k < hi; k += blockDim.x) {
    if (labels[k] != label) continue;
    const double b = bright[members[k]];
    if (b > best_b || (b == best_b && k > best_pos)) {
      best_b = b;
      best_pos = k;
    }
  }
  __shared__ double sb[128];
  __shared__ int sp[128];
  sb[threadIdx.x] = best_b;
  sp[threadIdx.x] = best_pos;
  __syncthreads();
  for (int o = 64; o > 0; o >>= 1) {
    if (threadIdx.x < o) {
      const double b = sb[threadIdx.x + o];
      const int pos = sp[threadIdx.x + o];
      if (b > sb[threadIdx.x] || (b == sb[threadIdx.x] && pos > sp[threadIdx.x])) {
        sb[threadIdx.x] = b;
        sp[threadIdx.x] = pos;
      }
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) survivor[region * max_labels + label] = sp[0] >= 0 ? members[sp[0]] : -1;
}

struct CandRegions {
  int count[CG_REGIONS];                       // members of every region, in the region's LIST order
  int member[CG_REGIONS][CG_MAX_MEMBERS];
  int combos[CG_REGIONS];                      // C(count, 5)
  int block_off[CG_REGIONS + 1];               // prefix of ceil(combos / CG_THREADS)
};

__device__ __forceinline__ int binom5(int n, int k) {   // C(n, k) for k <= 5, n <= 32
  if (k < 0 || n < k) return 0;
  long long r = 1;
  for (int i = 1; i <= k; ++i) r = r * (n - k + i) / i;
  return (int)r;
}

// rank -> the rank-th 5-combination of {0 .. n-1} in lexicographic order
__device__ __forceinline__ void unrank5(int rank, int n, int (&c)[5]) {
  int x = 0;
  for (int k = 0; k < 5; ++k) {
    // smallest x such that the combinations starting with x cover `rank`
    while (true) {
      const int with_x = binom5(n - x - 1, 4 - k);
      if (rank < with_x) break;
      rank -= with_x;
      ++x;
    }
    c[k] = x++;
  }
}

// The reference builds set(itertools.combinations(region list, 5) for every region) (cbpm:407-413): a
// set of TUPLES, so five stars listed by two regions in a different relative order count twice -- and
// that count is what its reduction loop compares with the maximum.  Measured, both tuples are the
// same polygon (its dict is keyed by hull order, cbpm:593).  Hence two answers per combination:
//   tuple duplicate: an EARLIER region lists all five stars in the same relative order  -> not counted
//   set duplicate:   an earlier region lists all five stars at all                      -> not written
template <bool WRITE>
__global__ void __launch_bounds__(CG_THREADS)
cand_kernel(CandRegions cr, int* __restrict__ block_tuples, int* __restrict__ block_sets, int* __restrict__ out) {
  __shared__ unsigned short member_mask[CG_MAX_STARS];        // star id -> regions that currently list it
  __shared__ unsigned char pos_in[CG_REGIONS][CG_MAX_STARS];  // position of the star in that region's list
  __shared__ int warp_cnt[2][CG_THREADS / 32];
  __shared__ int base_s;
  int region = 0;
  while (region + 1 < CG_REGIONS && (int)blockIdx.x >= cr.block_off[region + 1]) ++region;
  const int rank = ((int)blockIdx.x - cr.block_off[region]) * CG_THREADS + (int)threadIdx.x;
  for (int i = threadIdx.x; i < CG_MAX_STARS; i += CG_THREADS) member_mask[i] = 0;
  __syncthreads();
  if (threadIdx.x < CG_REGIONS)
    for (int k = 0; k < cr.count[threadIdx.x]; ++k) {
      const int id = cr.member[threadIdx.x][k];
      atomicOr(reinterpret_cast<unsigned*>(member_mask) + (id >> 1), (1u << threadIdx.x) << (16 * (id & 1)));
      pos_in[threadIdx.x][id] = (unsigned char)k;
    }
  __syncthreads();
  bool new_tuple = false, new_set = false;
  int id[5] = {0, 0, 0, 0, 0};
  if (rank < cr.combos[region]) {
    int c[5];
    unrank5(rank, cr.count[region], c);
    unsigned common = 0xFFFFu;
#pragma unroll
    for (int k = 0; k < 5; ++k) {
      id[k] = cr.member[region][c[k]];
      common &= member_mask[id[k]];
    }
    unsigned earlier = common & ((1u << region) - 1u);
    new_set = earlier == 0;
    new_tuple = true;
    while (earlier) {
      const int a = __ffs((int)earlier) - 1;
      earlier &= earlier - 1;
      bool same_order = true;
#pragma unroll
      for (int k = 0; k < 4; ++k) same_order = same_order && pos_in[a][id[k]] < pos_in[a][id[k + 1]];
      if (same_order) new_tuple = false;
    }
  }
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const unsigned ballot_t = __ballot_sync(0xffffffffu, new_tuple);
  const unsigned ballot_s = __ballot_sync(0xffffffffu, new_set);
  if (lane == 0) {
    warp_cnt[0][wid] = __popc(ballot_t);
    warp_cnt[1][wid] = __popc(ballot_s);
  }
  __syncthreads();
  int before = 0, total_t = 0, total_s = 0;
#pragma unroll
  for (int w = 0; w < CG_THREADS / 32; ++w) {
    before += (w < wid) ? warp_cnt[1][w] : 0;
    total_t += warp_cnt[0][w];
    total_s += warp_cnt[1][w];
  }
  if (!WRITE) {
    if (threadIdx.x == 0) {
      block_tuples[blockIdx.x] = total_t;
      block_sets[blockIdx.x] = total_s;
    }
    return;
  }
  if (threadIdx.x == 0) {
    int b = 0;
    for (int k = 0; k < (int)blockIdx.x; ++k) b += block_sets[k];   // <= a few hundred blocks
    base_s = b;
  }
  __syncthreads();
  if (new_set) {
    const int pos = base_s + before + __popc(ballot_s & ((1u << lane) - 1u));
#pragma unroll
    for (int k = 0; k < 5; ++k) out[pos * 5 + k] = id[k];
  }
}

// totals[0] = distinct tuples (the reference's len(set)), totals[1] = distinct star sets (table rows)
__global__ void cand_total_kernel(const int* __restrict__ block_tuples, const int* __restrict__ block_sets,
                                  int n_blocks, long long* __restrict__ totals) {
  long long t = 0, u = 0;
  for (int k = threadIdx.x; k < n_blocks; k += 32) {
    t += block_tuples[k];
    u += block_sets[k];
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    t += __shfl_xor_sync(0xffffffffu, t, o);
    u += __shfl_xor_sync(0xffffffffu, u, o);
  }
  if (threadIdx.x == 0) {
    totals[0] = t;
    totals[1] = u;
  }
}

}  // namespace ssp

extern "C" {

int ssp_star_regions(ssp_ctx* ctx, int image, uint16_t* region_mask, int64_t capacity) {
  using namespace ssp;
  if (!ctx) return SSP_ERR_INVALID;
  if (int rc = use_device(ctx)) return rc;
  SSP_REQUIRE(ctx, !ctx->star_off.empty() && image >= 0 && image < (int)ctx->star_off.size() - 1,
              SSP_ERR_STATE, "ssp_star_regions: image %d is not in the star bank", image);
  const long long s0 = ctx->star_off[(size_t)image];
  const int n = (int)(ctx->star_off[(size_t)image + 1] - s0);
  SSP_REQUIRE(ctx, capacity >= n, SSP_ERR_CAPACITY, "ssp_star_regions: capacity %lld < %d stars",
              (long long)capacity, n);
  if (n == 0) return SSP_OK;
  SSP_REQUIRE(ctx, region_mask != nullptr, SSP_ERR_INVALID, "ssp_star_regions: NULL output");
  NvtxRange range("candidate regions");
  SSP_CUDA_TRY(ctx, ctx->c_mask.ensure((size_t)n * 2 + 16));
  star_regions_kernel<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(
      ctx->s_xy.as<double>() + 2 * s0, n, (double)ctx->img_w, (double)ctx->img_h,
      ctx->c_mask.as<unsigned short>());
  SSP_CUDA_TRY(ctx, cudaGetLastError());
  SSP_CUDA_TRY(ctx, cudaMemcpyAsync(region_mask, ctx->c_mask.p, (size_t)n * 2, cudaMemcpyDeviceToHost,
                                    ctx->stream));
  SSP_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  return SSP_OK;
}

int ssp_cluster_brightest(ssp_ctx* ctx, int image, const int32_t* members, const int32_t* labels,
                          const int32_t* region_offsets, int n_regions, int max_labels,
                          int32_t* survivors) {
  using namespace ssp;
  if (!ctx) return SSP_ERR_INVALID;
  if (int rc = use_device(ctx)) return rc;
  SSP_REQUIRE(ctx, !ctx->star_off.empty() && image >= 0 && image < (int)ctx->star_off.size() - 1,
              SSP_ERR_STATE, "ssp_cluster_brightest: image %d is not in the star bank", image);
  SSP_REQUIRE(ctx, n_regions > 0 && n_regions <= 4096 && max_labels > 0 && max_labels <= 4096 &&
                       region_offsets != nullptr && survivors != nullptr,
              SSP_ERR_INVALID, "ssp_cluster_brightest: bad arguments");
  const int total = region_offsets[n_regions];
  SSP_REQUIRE(ctx, total >= 0 && (total == 0 || (members != nullptr && labels != nullptr)), SSP_ERR_INVALID,
              "ssp_cluster_brightest: bad member list");
  const long long s0 = ctx->star_off[(size_t)image];
  const int n_stars = (int)(ctx->star_off[(size_t)image + 1] - s0);
  for (int k = 0; k < total; ++k)
    SSP_REQUIRE(ctx, members[k] >= 0 && members[k] < n_stars && labels[k] >= 0 && labels[k] < max_labels,
                SSP_ERR_INVALID, "ssp_cluster_brightest: member %d out of range", k);
  NvtxRange range("candidate cluster selection");
  const size_t n_out = (size_t)n_regions * max_labels;
  SSP_CUDA_TRY(ctx, ctx->c_work.ensure(((size_t)2 * total + n_regions + 1 + n_out) * 4 + 64));
  int* d_members = ctx->c_work.as<int>();
  int* d_labels = d_members + total;
  int* d_off = d_labels + total;
  int* d_out = d_off + n_regions + 1;
  if (total > 0) {
    SSP_CUDA_TRY(ctx, cudaMemcpyAsync(d_members, members, (size_t)total * 4, cudaMemcpyHostToDevice, ctx->stream));
    SSP_CUDA_TRY(ctx, cudaMemcpyAsync(d_labels, labels, (size_t)total * 4, cudaMemcpyHostToDevice, ctx->stream));
  }
  SSP_CUDA_TRY(ctx, cudaMemcpyAsync(d_off, region_offsets, (size_t)(n_regions + 1) * 4, cudaMemcpyHostToDevice,
                                    ctx->stream));
  cluster_brightest_kernel<<<dim3((unsigned)n_regions, (unsigned)max_labels), 128, 0, ctx->stream>>>(
      d_members, d_labels, d_off, ctx->s_bright.as<double>() + s0, max_labels, d_out);
  SSP_CUDA_TRY(ctx, cudaGetLastError());
  SSP_CUDA_TRY(ctx, cudaMemcpyAsync(survivors, d_out, n_out * 4, cudaMemcpyDeviceToHost, ctx->stream));
  SSP_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  return SSP_OK;
}

int ssp_polygon_candidates(ssp_ctx* ctx, const int32_t* region_members, const int32_t* region_counts,
                           int n_regions, int64_t max_candidates, int64_t* n_tuples, int64_t* n_unique,
                           const int32_t** device_candidates) {
  using namespace ssp;
  if (!ctx) return SSP_ERR_INVALID;
  if (int rc = use_device(ctx)) return rc;
  SSP_REQUIRE(ctx, n_regions == CG_REGIONS, SSP_ERR_INVALID,
              "ssp_polygon_candidates: the reference has %d regions (cbpm:328-375), got %d", CG_REGIONS, n_regions);
  SSP_REQUIRE(ctx, region_members != nullptr && region_counts != nullptr && n_tuples != nullptr && n_unique != nullptr,
              SSP_ERR_INVALID, "ssp_polygon_candidates: NULL argument");
  CandRegions cr;
  int cursor = 0, blocks = 0;
  for (int r = 0; r < CG_REGIONS; ++r) {
    const int n = region_counts[r];
    SSP_REQUIRE(ctx, n >= 0 && n <= CG_MAX_MEMBERS, SSP_ERR_INVALID,
                "ssp_polygon_candidates: region %d lists %d stars, limit %d", r, n, CG_MAX_MEMBERS);
    cr.count[r] = n;
    for (int k = 0; k < n; ++k) {
      const int id = region_members[cursor + k];
      SSP_REQUIRE(ctx, id >= 0 && id < CG_MAX_STARS, SSP_ERR_INVALID,
                  "ssp_polygon_candidates: region %d: star id %d outside [0, %d)", r, id, CG_MAX_STARS);
      for (int j = 0; j < k; ++j)
        SSP_REQUIRE(ctx, region_members[cursor + j] != id, SSP_ERR_INVALID,
                    "ssp_polygon_candidates: region %d lists star %d twice", r, id);
      cr.member[r][k] = id;
    }
    for (int k = n; k < CG_MAX_MEMBERS; ++k) cr.member[r][k] = 0;
    cursor += n;
    long long c = 1;
    for (int i = 1; i <= 5; ++i) c = c * (n - 5 + i) / i;
    cr.combos[r] = n >= 5 ? (int)c : 0;
    cr.block_off[r] = blocks;
    blocks += (cr.combos[r] + CG_THREADS - 1) / CG_THREADS;
  }
  cr.block_off[CG_REGIONS] = blocks;
  *n_tuples = 0;
  *n_unique = 0;
  if (device_candidates) *device_candidates = nullptr;
  if (blocks == 0) return SSP_OK;
  NvtxRange range("candidate enumeration");
  SSP_CUDA_TRY(ctx, ctx->c_blocks.ensure((size_t)blocks * 8 + 32));
  int* d_tuples = ctx->c_blocks.as<int>();
  int* d_sets = d_tuples + blocks;
  long long* d_totals = reinterpret_cast<long long*>(ctx->c_blocks.as<char>() + ((size_t)blocks * 8 + 15) / 16 * 16);
  cand_kernel<false><<<(unsigned)blocks, CG_THREADS, 0, ctx->stream>>>(cr, d_tuples, d_sets, nullptr);
  SSP_CUDA_TRY(ctx, cudaGetLastError());
  cand_total_kernel<<<1, 32, 0, ctx->stream>>>(d_tuples, d_sets, blocks, d_totals);
  SSP_CUDA_TRY(ctx, cudaGetLastError());
  long long totals[2] = {0, 0};
  SSP_CUDA_TRY(ctx, cudaMemcpyAsync(totals, d_totals, 16, cudaMemcpyDeviceToHost, ctx->stream));
  SSP_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  *n_tuples = totals[0];
  *n_unique = totals[1];
  // the reference's loop condition looks at the TUPLE count (cbpm:383); the host retries with fewer stars
  if (totals[0] == 0 || totals[0] > max_candidates || !device_candidates) return SSP_OK;
  SSP_CUDA_TRY(ctx, ctx->c_cand.ensure((size_t)totals[1] * 20 + 16));
  cand_kernel<true><<<(unsigned)blocks, CG_THREADS, 0, ctx->stream>>>(cr, d_tuples, d_sets, ctx->c_cand.as<int>());
  SSP_CUDA_TRY(ctx, cudaGetLastError());
  *device_candidates = ctx->c_cand.as<int>();
  return SSP_OK;
}

}  // extern "C"
