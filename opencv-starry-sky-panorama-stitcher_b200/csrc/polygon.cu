// Kernel 2 (star -> sphere projection, polygon measure) and kernel 3 (polygon vs polygon) of
// libssp_b200.so.  FP64 throughout and in the reference's operation order
// (custom_bf_polygon_matcher.py:70-135, :500-600, :722-740, :992-1067): the spherical law of
// cosines used there is ill-conditioned, FP32 cannot hold the 1e-5 rad parity bound.
#include "common.cuh"

namespace ssp {

// ---------------------------------------------------------------------------------------------
// star -> unit ray (angle_distance, cbpm:101-106, unit_vector cbpm:74-76)
// ---------------------------------------------------------------------------------------------
__global__ void star_rays_kernel(const double* __restrict__ xy, long long n, double half_w,
                                 double half_h, double focal, double* __restrict__ rays) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double vx = xy[2 * i] - half_w;
  const double vy = xy[2 * i + 1] - half_h;
  const double vz = focal;
  const double nn = __dadd_rn(__dadd_rn(__dmul_rn(vx, vx), __dmul_rn(vy, vy)), __dmul_rn(vz, vz));
  const double norm = sqrt(nn);
  rays[3 * i] = __ddiv_rn(vx, norm);
  rays[3 * i + 1] = __ddiv_rn(vy, norm);
  rays[3 * i + 2] = __ddiv_rn(vz, norm);
}

struct StarView {
  const double* xy;      // n x 2
  const double* bright;  // n
  const double* rays;    // n x 3
  int n;
};

// angle_between of two unit rays: arccos(clip(dot, -1, 1))  (cbpm:78-84)
__device__ __forceinline__ double ray_angle(const double* a, const double* b) {
  double d = __dadd_rn(__dadd_rn(__dmul_rn(a[0], b[0]), __dmul_rn(a[1], b[1])), __dmul_rn(a[2], b[2]));
  d = fmin(fmax(d, -1.0), 1.0);
  return acos(d);
}

// sign-exact 2-D orientation for float32-valued pixel coordinates (products are exact in
// double); > 0: c is left of a->b in (x right, y up) axes.
__device__ __forceinline__ double orient2d(double ax, double ay, double bx, double by, double cx,
                                           double cy) {
  return __dsub_rn(__dmul_rn(bx - ax, cy - ay), __dmul_rn(by - ay, cx - ax));
}

// One thread per candidate polygon: the body of get_polygons_characteristics (cbpm:500-600)
// plus the normalisation of cbpm:722-740.  Results go to per-candidate slots; a stable
// compaction keeps the valid ones afterwards.
__device__ __forceinline__ void
poly_measure_one(const StarView& sv, const int* __restrict__ cand, long long pidx, double min_side,
                 unsigned char* __restrict__ status, int* __restrict__ flags,
                 int* __restrict__ n_domain, int* __restrict__ o_idx,
                 double* __restrict__ o_angle, double* __restrict__ o_side,
                 double* __restrict__ o_bright, double* __restrict__ o_feat) {
  int id[5];
  double x[5], y[5];
  bool bad = false;
#pragma unroll
  for (int k = 0; k < 5; ++k) {
    id[k] = cand[pidx * 5 + k];
    if (id[k] < 0 || id[k] >= sv.n) {
      bad = true;
      id[k] = 0;
    }
    x[k] = sv.xy[2 * (long long)id[k]];
    y[k] = sv.xy[2 * (long long)id[k] + 1];
  }
  if (bad) {
    status[pidx] = SSP_POLY_BAD_INDEX;
    flags[pidx] = 0;
    return;
  }
  // (1) reject polygons with a pixel distance below min_side (cbpm:515, dist_between_2_points)
  bool short_side = false;
#pragma unroll
  for (int i = 0; i < 5; ++i)
#pragma unroll
    for (int j = i + 1; j < 5; ++j) {
      const double dx = x[j] - x[i], dy = y[j] - y[i];
      const double d = sqrt(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)));
      short_side |= (d < min_side);
    }
  if (short_side) {
    status[pidx] = SSP_POLY_SHORT_SIDE;
    flags[pidx] = 0;
    return;
  }
  // (2) convex hull of the 5 points by gift wrapping; all 5 must be strict hull vertices
  //     (alphashape(pts, 0) == MultiPoint.convex_hull drops interior and collinear points,
  //     cbpm:520-526).  Start at the lowest point (min y, then min x).
  int start = 0;
#pragma unroll
  for (int k = 1; k < 5; ++k)
    if (y[k] < y[start] || (y[k] == y[start] && x[k] < x[start])) start = k;
  int ord[5];
  int cnt = 1, cur = start;
  ord[0] = start;
  bool closed = false;
  for (int step = 0; step < 5; ++step) {
    int nxt = -1;
    for (int s = 0; s < 5; ++s) {
      if (s == cur) continue;
      if (nxt < 0) {
        nxt = s;
        continue;
      }
      const double o = orient2d(x[cur], y[cur], x[nxt], y[nxt], x[s], y[s]);
      if (o < 0.0) {
        nxt = s;
      } else if (o == 0.0) {
        const double dn = (x[nxt] - x[cur]) * (x[nxt] - x[cur]) + (y[nxt] - y[cur]) * (y[nxt] - y[cur]);
        const double ds = (x[s] - x[cur]) * (x[s] - x[cur]) + (y[s] - y[cur]) * (y[s] - y[cur]);
        if (ds > dn) nxt = s;
      }
    }
    cur = nxt;
    if (cur == start) {
      closed = true;
      break;
    }
    if (cnt < 5) ord[cnt] = cur;
    ++cnt;
  }
  if (!closed || cnt != 5) {
    status[pidx] = SSP_POLY_NOT_CONVEX;
    flags[pidx] = 0;
    return;
  }
  // gift wrapping above walks counter-clockwise in (x right, y up) axes, i.e. positive
  // shoelace area; the reference keeps the ring with NEGATIVE signed area (cbpm:537-543):
  // same start vertex, opposite direction.
  int ring[5];
  ring[0] = ord[0];
  ring[1] = ord[4];
  ring[2] = ord[3];
  ring[3] = ord[2];
  ring[4] = ord[1];

  // (3) spherical measures.  side[i] = angle(ring[i], ring[i+1]) (cbpm:569-576);
  //     vertex angle i from (prev, cur, next) by the spherical law of cosines (cbpm:127-135)
  double ray[5][3];
#pragma unroll
  for (int k = 0; k < 5; ++k) {
    const long long s = id[ring[k]];
    ray[k][0] = sv.rays[3 * s];
    ray[k][1] = sv.rays[3 * s + 1];
    ray[k][2] = sv.rays[3 * s + 2];
  }
  double side[5], angle[5];
#pragma unroll
  for (int k = 0; k < 5; ++k) side[k] = ray_angle(ray[k], ray[(k + 1) % 5]);
  bool domain = false;
#pragma unroll
  for (int k = 0; k < 5; ++k) {
    const int prev = (k + 4) % 5, next = (k + 1) % 5;
    const double b = side[prev];                   // angle_p1_p2
    const double c = side[k];                      // angle_p2_p3
    const double a = ray_angle(ray[next], ray[prev]);  // angle_p3_p1
    const double num = __dsub_rn(cos(a), __dmul_rn(cos(b), cos(c)));
    const double den = __dmul_rn(sin(b), sin(c));
    const double v = __ddiv_rn(num, den);
    if (!(fabs(v) <= 1.0)) domain = true;  // math.acos raises ValueError (cbpm:134)
    angle[k] = acos(v);
  }
  if (domain) {
    status[pidx] = SSP_POLY_DOMAIN;
    flags[pidx] = 0;
    atomicAdd(n_domain, 1);
    return;
  }
  // (4) rotate so the smallest angle (first occurrence) is element 0 (cbpm:578-591)
  int rot = 0;
#pragma unroll
  for (int k = 1; k < 5; ++k)
    if (angle[k] < angle[rot]) rot = k;
  double max_side = 0.0, max_bright = 0.0;
  double r_angle[5], r_side[5], r_bright[5];
  int r_idx[5];
#pragma unroll
  for (int k = 0; k < 5; ++k) {
    const int src = (k + rot) % 5;
    r_idx[k] = id[ring[src]];
    r_angle[k] = angle[src];
    r_side[k] = side[src];
    r_bright[k] = sv.bright[r_idx[k]];
    max_side = fmax(max_side, r_side[k]);
    max_bright = fmax(max_bright, r_bright[k]);
  }
  // (5) normalise: angles / pi, sides / max side, brightness / max brightness (cbpm:722-740)
  const double pi = 3.141592653589793;
#pragma unroll
  for (int k = 0; k < 5; ++k) {
    o_idx[pidx * 5 + k] = r_idx[k];
    o_angle[pidx * 5 + k] = r_angle[k];
    o_side[pidx * 5 + k] = r_side[k];
    o_bright[pidx * 5 + k] = r_bright[k];
    o_feat[pidx * 15 + k] = __ddiv_rn(r_angle[k], pi);
    o_feat[pidx * 15 + 5 + k] = __ddiv_rn(r_side[k], max_side);
    o_feat[pidx * 15 + 10 + k] = __ddiv_rn(r_bright[k], max_bright);
  }
  status[pidx] = SSP_POLY_VALID;
  flags[pidx] = 1;
}

__global__ void __launch_bounds__(128)
poly_measure_kernel(StarView sv, const int* __restrict__ cand, long long n_cand, double min_side,
                    unsigned char* __restrict__ status, int* __restrict__ flags,
                    int* __restrict__ n_domain, int* __restrict__ o_idx,
                    double* __restrict__ o_angle, double* __restrict__ o_side,
                    double* __restrict__ o_bright, double* __restrict__ o_feat) {
  const long long pidx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (pidx >= n_cand) return;
  poly_measure_one(sv, cand, pidx, min_side, status, flags, n_domain, o_idx, o_angle, o_side, o_bright, o_feat);
}

// Several images in ONE launch (ssp_polygon_measure_batch): the candidate tables sit back to back,
// every image owns whole 128-thread blocks.  A C4 step measured 21 images with 21 launches of ~118
// blocks each, ~25 us apiece and latency-bound on 148 SMs; together they are one grid of ~2 500 blocks.
struct MeasureImage {
  StarView sv;
  long long cand_off;    // first candidate of the image in the concatenated table (= scratch index)
  int n;                 // candidates of the image
  int block_off;         // first block of the image
  int* o_idx;            // the image's polygon bank
  double *o_angle, *o_side, *o_bright, *o_feat;
  long long* slot;       // {count, n_domain} of the image (device memory, copied to the pinned slot afterwards)
};

__device__ __forceinline__ int measure_image_of_block(const MeasureImage* __restrict__ imgs, int n_images, int block) {
  int lo = 0, hi = n_images - 1;
  while (lo < hi) {
    const int mid = (lo + hi + 1) >> 1;
    if (imgs[mid].block_off <= block) lo = mid; else hi = mid - 1;
  }
  return lo;
}

__global__ void __launch_bounds__(128)
poly_measure_batch_kernel(const MeasureImage* __restrict__ imgs, int n_images, const int* __restrict__ cand,
                          double min_side, unsigned char* __restrict__ status, int* __restrict__ flags,
                          int* __restrict__ o_idx, double* __restrict__ o_angle, double* __restrict__ o_side,
                          double* __restrict__ o_bright, double* __restrict__ o_feat) {
  const MeasureImage im = imgs[measure_image_of_block(imgs, n_images, (int)blockIdx.x)];
  const int local = ((int)blockIdx.x - im.block_off) * 128 + (int)threadIdx.x;
  if (local >= im.n) return;
  // the domain-error counter of the image is the second word of its slot
  poly_measure_one(im.sv, cand, im.cand_off + local, min_side, status, flags,
                   reinterpret_cast<int*>(im.slot + 1), o_idx, o_angle, o_side, o_bright, o_feat);
}

// Stable compaction of the valid candidates.  Every block derives its own output base from the
// flags in front of it (at most 15 000 ints, L2-resident) -- round 1 ran a separate single-CTA scan
// kernel between measure and scatter, 12 us per image, as much as half a measure launch.  The last
// block also leaves the total where the host (or the pinned no-wait slot) reads it.
__global__ void __launch_bounds__(128)
poly_scatter_kernel(const int* __restrict__ flags, long long* __restrict__ total_out,
                    long long n_cand, const int* __restrict__ t_idx,
                    const double* __restrict__ t_angle,
                    const double* __restrict__ t_side,
                    const double* __restrict__ t_bright,
                    const double* __restrict__ t_feat, int* __restrict__ o_idx,
                    double* __restrict__ o_angle, double* __restrict__ o_side,
                    double* __restrict__ o_bright, double* __restrict__ o_feat) {
  __shared__ int wsum[4];
  __shared__ int wcnt[4];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const long long first = (long long)blockIdx.x * blockDim.x;
  int before = 0;
  for (long long k = threadIdx.x; k < first; k += blockDim.x) before += flags[k];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) before += __shfl_xor_sync(0xffffffffu, before, o);
  const long long i = first + threadIdx.x;
  const bool f = i < n_cand && flags[i] != 0;
  const unsigned ballot = __ballot_sync(0xffffffffu, f);
  if (lane == 0) {
    wsum[wid] = before;
    wcnt[wid] = __popc(ballot);
  }
  __syncthreads();
  long long d = (long long)wsum[0] + wsum[1] + wsum[2] + wsum[3];
  int block_total = 0;
#pragma unroll
  for (int w = 0; w < 4; ++w) {
    d += (w < wid) ? wcnt[w] : 0;
    block_total += wcnt[w];
  }
  if (blockIdx.x == gridDim.x - 1 && threadIdx.x == 0)
    *total_out = (long long)wsum[0] + wsum[1] + wsum[2] + wsum[3] + block_total;
  if (!f) return;
  d += __popc(ballot & ((1u << lane) - 1u));
#pragma unroll
  for (int k = 0; k < 5; ++k) {
    o_idx[d * 5 + k] = t_idx[i * 5 + k];
    o_angle[d * 5 + k] = t_angle[i * 5 + k];
    o_side[d * 5 + k] = t_side[i * 5 + k];
    o_bright[d * 5 + k] = t_bright[i * 5 + k];
  }
#pragma unroll
  for (int k = 0; k < 15; ++k) o_feat[d * 15 + k] = t_feat[i * 15 + k];
}

__global__ void __launch_bounds__(128)
poly_scatter_batch_kernel(const MeasureImage* __restrict__ imgs, int n_images, const int* __restrict__ flags,
                          const int* __restrict__ t_idx, const double* __restrict__ t_angle,
                          const double* __restrict__ t_side, const double* __restrict__ t_bright,
                          const double* __restrict__ t_feat) {
  __shared__ int wsum[4];
  __shared__ int wcnt[4];
  const MeasureImage im = imgs[measure_image_of_block(imgs, n_images, (int)blockIdx.x)];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int first = ((int)blockIdx.x - im.block_off) * 128;     // inside the image
  const int* f = flags + im.cand_off;
  int before = 0;
  for (int k = threadIdx.x; k < first; k += 128) before += f[k];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) before += __shfl_xor_sync(0xffffffffu, before, o);
  const int local = first + (int)threadIdx.x;
  const bool keep = local < im.n && f[local] != 0;
  const unsigned ballot = __ballot_sync(0xffffffffu, keep);
  if (lane == 0) {
    wsum[wid] = before;
    wcnt[wid] = __popc(ballot);
  }
  __syncthreads();
  long long d = (long long)wsum[0] + wsum[1] + wsum[2] + wsum[3];
  int block_total = 0;
#pragma unroll
  for (int w = 0; w < 4; ++w) {
    d += (w < wid) ? wcnt[w] : 0;
    block_total += wcnt[w];
  }
  const bool last_block = first + 128 >= im.n;
  if (last_block && threadIdx.x == 0) im.slot[0] = (long long)wsum[0] + wsum[1] + wsum[2] + wsum[3] + block_total;
  if (!keep) return;
  d += __popc(ballot & ((1u << lane) - 1u));
  const long long i = im.cand_off + local;
#pragma unroll
  for (int k = 0; k < 5; ++k) {
    im.o_idx[d * 5 + k] = t_idx[i * 5 + k];
    im.o_angle[d * 5 + k] = t_angle[i * 5 + k];
    im.o_side[d * 5 + k] = t_side[i * 5 + k];
    im.o_bright[d * 5 + k] = t_bright[i * 5 + k];
  }
#pragma unroll
  for (int k = 0; k < 15; ++k) im.o_feat[d * 15 + k] = t_feat[i * 15 + k];
}

// ---------------------------------------------------------------------------------------------
// kernel 3: polygon versus polygon (cbpm:992-1067, :1145-1148)
// ---------------------------------------------------------------------------------------------
constexpr int PC_THREADS = 128;  // rows per CTA (one polygon of image 1 per thread)
constexpr int PC_TILE = 128;     // columns (polygons of image 2) per shared-memory tile

// Squared distance in numpy's float64 evaluation order: per property a sequential sum of the 5
// squared differences, then (angles + sides) + brightness; no FMA contraction so every
// intermediate rounds exactly like np.square / np.sum / the two additions of cbpm:1011-1016.
template <int SHIFT>
__device__ __forceinline__ double poly_dist2_f64(const double (&a)[15], const double* __restrict__ b) {
  double tot = 0.0;
#pragma unroll
  for (int g = 0; g < 3; ++g) {
    double acc = 0.0;
#pragma unroll
    for (int k = 0; k < 5; ++k) {
      const double d = __dsub_rn(a[g * 5 + k], b[g * 5 + (k + SHIFT) % 5]);
      const double s = __dmul_rn(d, d);
      acc = (k == 0) ? s : __dadd_rn(acc, s);
    }
    tot = (g == 0) ? acc : __dadd_rn(tot, acc);
  }
  return tot;
}

#ifndef POLY_G0
#define POLY_G0 1   // first and second property group examined by the pruning compare: sides, then
#define POLY_G1 0   // angles (tools/poly_order_sweep.sh: 4.0 ms at C4 against 4.9 for angles first, 7.0 unpruned)
#endif
// one property group (g = 0 angles, 1 sides, 2 brightnesses) of poly_dist2_f64<0>
template <int G>
__device__ __forceinline__ double poly_group_f64(const double (&a)[15], const double* __restrict__ b) {
  double acc = 0.0;
#pragma unroll
  for (int k = 0; k < 5; ++k) {
    const double d = __dsub_rn(a[G * 5 + k], b[G * 5 + k]);
    const double s = __dmul_rn(d, d);
    acc = (k == 0) ? s : __dadd_rn(acc, s);
  }
  return acc;
}

template <int SHIFT>
__device__ __forceinline__ float poly_dist2_f32(const float (&a)[15], const float* __restrict__ b) {
  float tot = 0.f;
#pragma unroll
  for (int g = 0; g < 3; ++g)
#pragma unroll
    for (int k = 0; k < 5; ++k) {
      const float d = a[g * 5 + k] - b[g * 5 + (k + SHIFT) % 5];
      tot = fmaf(d, d, tot);
    }
  return tot;
}

// FP64 parity kernel.  Row argmin is taken on sqrt(dist2) exactly as np.argmin sees it: two
// different dist2 values can round to the same square root, and the reference then keeps the
// earlier column.  `thr` is the dist2 of the last examined improvement, so the sqrt is only
// evaluated O(log P2) times per row.
template <bool MIN_SHIFTS>
__global__ void __launch_bounds__(PC_THREADS)
poly_compare_f64_kernel(const PolyPair* __restrict__ pairs, int* __restrict__ argmin,
                        double* __restrict__ mindist) {
  __shared__ __align__(16) double tile[PC_TILE * 15];
  const PolyPair pp = pairs[blockIdx.x];
  const int row0 = blockIdx.y * PC_THREADS;
  if (row0 >= pp.p1) return;
  const int r = row0 + threadIdx.x;
  const bool live = r < pp.p1;
  double a[15];
#pragma unroll
  for (int k = 0; k < 15; ++k) a[k] = live ? pp.f1[(long long)r * 15 + k] : 0.0;
  double thr = INFINITY, best = INFINITY;
  int best_c = 0;
  for (int c0 = 0; c0 < pp.p2; c0 += PC_TILE) {
    const int ncol = min(PC_TILE, pp.p2 - c0);
    __syncthreads();
    for (int i = threadIdx.x; i < ncol * 15; i += PC_THREADS)
      tile[i] = pp.f2[(long long)c0 * 15 + i];
    __syncthreads();
    if (live) {
      for (int j = 0; j < ncol; ++j) {
        const double* b = tile + j * 15;
        double v;
        if (MIN_SHIFTS) {
          v = poly_dist2_f64<0>(a, b);
          v = fmin(v, poly_dist2_f64<1>(a, b));
          v = fmin(v, poly_dist2_f64<2>(a, b));
          v = fmin(v, poly_dist2_f64<3>(a, b));
          v = fmin(v, poly_dist2_f64<4>(a, b));
        } else {
          // Exact pruning: every term is >= 0 and floating-point addition of a non-negative
          // number never decreases the sum, so a partial sum that already reaches the row's
          // current best rules the column out; the values that ARE completed go through
          // exactly the operations of poly_dist2_f64<0>.  (The branch is taken per warp: the
          // angle group alone rules a column out for all 32 rows about a quarter of the time,
          // angles + sides nearly always.)
          // Any group, and any sum of two groups, is such a lower bound, so the groups are
          // examined in the order that prunes best (POLY_G0, POLY_G1, then the third); the
          // completed value is still summed (angles + sides) + brightnesses.
          double g[3];
          g[POLY_G0] = poly_group_f64<POLY_G0>(a, b);
          if (!(g[POLY_G0] < thr)) continue;
          g[POLY_G1] = poly_group_f64<POLY_G1>(a, b);
          if (!(__dadd_rn(g[POLY_G0], g[POLY_G1]) < thr)) continue;
          g[3 - POLY_G0 - POLY_G1] = poly_group_f64<3 - POLY_G0 - POLY_G1>(a, b);
          v = __dadd_rn(__dadd_rn(g[0], g[1]), g[2]);
        }
        if (v < thr) {
          const double s = sqrt(v);
          if (s < best) {
            best = s;
            best_c = c0 + j;
          }
          thr = v;
        }
      }
    }
  }
  if (live) {
    argmin[pp.out_row + r] = best_c;
    mindist[pp.out_row + r] = best;
  }
}

// FP64 parity kernel, shift 0, with COMPACTED pruning (round 2).  In the kernel above every lane
// prunes against its own row's threshold, so after the first property group (which all 32 lanes
// evaluate for every column) the warp drags 2-3 surviving lanes through the other two groups for
// ~40 % of the columns: ncu showed 18.6 of 32 lanes active per FP64 instruction.  Here the first
// group is only a thread-per-row FILTER, and a survivor is only QUEUED: (row lane, column) goes to a
// per-warp queue in shared memory (ballot + prefix), and whenever 32 entries are waiting the warp
// evaluates them as one dense batch -- lane i takes entry i, whatever row it belongs to (the rows'
// features sit in shared memory too) -- from scratch, in exactly the operations and the order of
// poly_dist2_f64<0>.  An entry that beats its row's threshold is handed to the lane that owns the row
// (shuffle broadcast, ascending lane = ascending column order per row, rare: O(log P2) per row),
// which applies the sequential rule of the reference unchanged.  Thresholds used for pruning are the
// row's own, possibly a few entries stale, i.e. never too small: exact.
//
// The filter only has to be SAFE (never drop a column whose exact first-group sum is below the row's
// threshold), not bit-equal, and ncu showed what it has to be cheap in: the shared-memory data pipe
// (l1tex__data_pipe_lsu_wavefronts 92 % of peak -- a third of it the filter's broadcast loads, one
// wavefront per 8 bytes per column) and instruction issue (45 per column and warp, 72 % of peak).
//   PQ_FILTER 2 (default): packed single precision -- the first group's features as floats, k-major,
//     four columns per LDS.128, two columns per FADD2 / FFMA2: half the wavefronts, half the arithmetic
//     instructions of the filter, none of them on the FP64 pipe.  Safety: with u = 2^-24, F >= every
//     abs(feature) involved, delta_k the real differences and e_k the float ones,
//     abs(e_k - delta_k) <= 4.0000003 u F, so sum e_k^2 <= sum delta_k^2 + 80.0001 u F^2, and the float
//     accumulation (five roundings of non-negative terms, FMA) gives S_f <= (sum e_k^2)(1 + u)^5.  A
//     column is dropped iff S_f >= T_f with T_f >= (thr (1 + 2^-48) + 81 u F^2)(1 + 6 u), which implies
//     sum delta_k^2 >= thr (1 + 2^-48) and so the FP64 sum (>= (1 - 2^-50) sum delta_k^2) >= thr.  F is
//     taken per row and per column tile; where F^2 could overflow a float, T_f is NaN and nothing drops.
//   PQ_FILTER 1: FP64 with fused multiply-adds (10 instead of 14 FP64 instructions per column): with the
//     same rounded differences the fused sum Gf and numpy's G both lie within (1 +- u)^5 of the real
//     sum, so a column dropped at Gf >= thr (1 + 2^-40) has G >= thr.
#ifndef PQ_TILE_COLS
#define PQ_TILE_COLS 64
#endif
#ifndef PQ_UNROLL
#define PQ_UNROLL 8      // columns whose first-group sums are in flight together (independent chains)
#endif
#ifndef PQ_FILTER
#define PQ_FILTER 2
#endif
#ifndef PQ_MIN_BLOCKS
#define PQ_MIN_BLOCKS 7  // resident blocks per SM the register allocation is held to (72 registers, no spills)
#endif
static_assert(PQ_UNROLL % 4 == 0 && PQ_TILE_COLS % PQ_UNROLL == 0, "column steps are whole groups of four");
constexpr int PQ_TILE = PQ_TILE_COLS;        // columns per shared-memory tile
constexpr int PQ_QUEUE = 32 * (PQ_UNROLL + 1);   // queue entries per warp: 31 left over + one column step of PQ_UNROLL x 32
constexpr int PQ_G2 = 3 - POLY_G0 - POLY_G1; // the third property group
constexpr int PQ_RS = 15;                    // values per row in shared memory: groups POLY_G1, PQ_G2, POLY_G0 (odd: see rows_lo)
constexpr int PQ_TFS = PQ_TILE + 4;          // floats per feature of the float tile (16-byte aligned rows)
constexpr double PQ_LOOSE = 1.0 + 9.094947017729282e-13;   // 1 + 2^-40

// Threshold of the float filter for the FP64 threshold thr and the margin 81 u F^2 (see above).
__device__ __forceinline__ float pq_float_threshold(double thr, double margin, bool f_ok) {
  const double t = (thr * (1.0 + 3.552713678800501e-15) + margin + 1e-36) * (1.0 + 6.0 * 5.9604644775390625e-8);
  return f_ok ? __double2float_ru(t) : __int_as_float(0x7fc00000);
}

__global__ void __launch_bounds__(PC_THREADS, PQ_MIN_BLOCKS)
poly_compare_f64_queue_kernel(const PolyPair* __restrict__ pairs, int* __restrict__ argmin,
                              double* __restrict__ mindist, unsigned long long* __restrict__ executed_flops) {
  __shared__ __align__(16) double tile[(PQ_TILE + PQ_UNROLL) * 15];
  // rows and thresholds as separate low / high words: a batch reads them at 32 arbitrary rows, and a
  // 32-bit access at an odd row stride has no bank conflict whatever the rows are (an LDS.64 there took
  // 3.9 wavefronts instead of 2: rows r and r + 16 share a bank pair)
  __shared__ unsigned rows_lo[PC_THREADS * PQ_RS], rows_hi[PC_THREADS * PQ_RS];
  __shared__ unsigned thr_lo[PC_THREADS], thr_hi[PC_THREADS];
  __shared__ int q_rc[PC_THREADS / 32][PQ_QUEUE];             // row lane << 16 | column in tile
#if PQ_FILTER == 2
  __shared__ __align__(16) float tile_f[5 * PQ_TFS];           // first group of the tile's columns, feature-major
  __shared__ unsigned tile_fmax[2];                            // max abs first-group feature of the tile (float bits)
#endif
  const PolyPair pp = pairs[blockIdx.x];
  const int row0 = blockIdx.y * PC_THREADS;
  if (row0 >= pp.p1) return;
  const int r = row0 + threadIdx.x;
  const bool live = r < pp.p1;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#if PQ_FILTER == 2
  float2 na[5];          // (-a_k, -a_k) of the first group
  float f_row = 0.0f;    // >= abs of the row's first-group features
#else
  double a0[5];
#endif
#pragma unroll
  for (int k = 0; k < 5; ++k) {
    const double g0v = live ? pp.f1[(long long)r * 15 + POLY_G0 * 5 + k] : 0.0;
    const double g1v = live ? pp.f1[(long long)r * 15 + POLY_G1 * 5 + k] : 0.0;
    const double g2v = live ? pp.f1[(long long)r * 15 + PQ_G2 * 5 + k] : 0.0;
    rows_lo[threadIdx.x * PQ_RS + k] = (unsigned)__double2loint(g1v);
    rows_hi[threadIdx.x * PQ_RS + k] = (unsigned)__double2hiint(g1v);
    rows_lo[threadIdx.x * PQ_RS + 5 + k] = (unsigned)__double2loint(g2v);
    rows_hi[threadIdx.x * PQ_RS + 5 + k] = (unsigned)__double2hiint(g2v);
    rows_lo[threadIdx.x * PQ_RS + 10 + k] = (unsigned)__double2loint(g0v);
    rows_hi[threadIdx.x * PQ_RS + 10 + k] = (unsigned)__double2hiint(g0v);
#if PQ_FILTER == 2
    const float af = __double2float_rn(g0v);
    na[k] = make_float2(-af, -af);
    f_row = fmaxf(f_row, __double2float_ru(fabs(g0v)));
#else
    a0[k] = g0v;
#endif
  }
  // a dead lane (row beyond p1) has threshold -1: nothing it queues ever qualifies
  double thr = live ? INFINITY : -1.0, best = INFINITY;
#if PQ_FILTER == 2
  float thr_filter = 0.0f;                       // set per tile
  double margin = 0.0;
  bool f_ok = true;
  if (threadIdx.x < 2) tile_fmax[threadIdx.x] = 0u;
#else
  double thr_filter = thr;                       // what the first-group filter compares with (thr, loosened)
#endif
  int best_c = 0;
  thr_lo[threadIdx.x] = (unsigned)__double2loint(thr);
  thr_hi[threadIdx.x] = (unsigned)__double2hiint(thr);
  unsigned flops = 0, flops_f32 = 0;   // executed by this thread: FP64 (batches, FP64 filter), single precision (filter)
  int q_n = 0;                                   // entries waiting in this warp's queue (warp-uniform)
  int* my_rc = q_rc[wid];
  const unsigned* my_rows_lo = rows_lo + wid * 32 * PQ_RS;
  const unsigned* my_rows_hi = rows_hi + wid * 32 * PQ_RS;
  unsigned* my_thr_lo = thr_lo + wid * 32;
  unsigned* my_thr_hi = thr_hi + wid * 32;

  // evaluates up to 32 queued entries [first, first + n) as one dense batch
  auto run_batch = [&](int first, int n, int c0, int ncol) {
    double v = 0.0;
    int rl = 0, col = 0;
    bool cand = false;
    if (lane < n) {
      const int rc = my_rc[first + lane];
      rl = rc >> 16;
      col = rc & 0xFFFF;
    }
    if (lane < n && col < ncol) {   // a padding column of a ragged tile may get here when nothing is dropped
      const double* b = tile + col * 15;
      const unsigned* alo = my_rows_lo + rl * PQ_RS;
      const unsigned* ahi = my_rows_hi + rl * PQ_RS;
      auto ar = [&](int k) { return __hiloint2double((int)ahi[k], (int)alo[k]); };
      const double t = __hiloint2double((int)my_thr_hi[rl], (int)my_thr_lo[rl]);
      double g0 = 0.0;     // the filter's sum was only a bound: the real first group, numpy's order
#pragma unroll
      for (int k = 0; k < 5; ++k) {
        const double d = __dsub_rn(ar(10 + k), b[POLY_G0 * 5 + k]);
        const double sq = __dmul_rn(d, d);
        g0 = (k == 0) ? sq : __dadd_rn(g0, sq);
      }
      double g1 = 0.0;
#pragma unroll
      for (int k = 0; k < 5; ++k) {
        const double d = __dsub_rn(ar(k), b[POLY_G1 * 5 + k]);
        const double sq = __dmul_rn(d, d);
        g1 = (k == 0) ? sq : __dadd_rn(g1, sq);
      }
      flops += 14 + 15;
      if (g0 < t && __dadd_rn(g0, g1) < t) {
        double g2 = 0.0;
#pragma unroll
        for (int k = 0; k < 5; ++k) {
          const double d = __dsub_rn(ar(5 + k), b[PQ_G2 * 5 + k]);
          const double sq = __dmul_rn(d, d);
          g2 = (k == 0) ? sq : __dadd_rn(g2, sq);
        }
        flops += 16;
        double g[3];
        g[POLY_G0] = g0;
        g[POLY_G1] = g1;
        g[PQ_G2] = g2;
        v = __dadd_rn(__dadd_rn(g[0], g[1]), g[2]);   // (angles + sides) + brightnesses, cbpm:1011-1016
        cand = v < t;
      }
    }
    // improvements go to the lane that owns the row, in ascending entry (= column) order
    unsigned todo = __ballot_sync(0xffffffffu, cand);
    while (todo) {
      const int src = __ffs((int)todo) - 1;
      todo &= todo - 1;
      const int o_rl = __shfl_sync(0xffffffffu, rl, src);
      const int o_col = __shfl_sync(0xffffffffu, col, src);
      const double o_v = __shfl_sync(0xffffffffu, v, src);
      if (lane == o_rl && o_v < thr) {
        const double sv = sqrt(o_v);
        if (sv < best) {
          best = sv;
          best_c = c0 + o_col;
        }
        thr = o_v;
#if PQ_FILTER == 2
        thr_filter = pq_float_threshold(o_v, margin, f_ok);
#else
        thr_filter = o_v * PQ_LOOSE;
#endif
        my_thr_lo[lane] = (unsigned)__double2loint(o_v);
        my_thr_hi[lane] = (unsigned)__double2hiint(o_v);
      }
      __syncwarp();
    }
  };

  for (int c0 = 0, it = 0; c0 < pp.p2; c0 += PQ_TILE, ++it) {
    const int ncol = min(PQ_TILE, pp.p2 - c0);
    const int ncol_pad = ((ncol + PQ_UNROLL - 1) / PQ_UNROLL) * PQ_UNROLL;
    __syncthreads();
#if PQ_FILTER == 2
    float fm = 0.0f;
#endif
    for (int i = threadIdx.x; i < ncol * 15; i += PC_THREADS) {
      const double x = pp.f2[(long long)c0 * 15 + i];
      tile[i] = x;
#if PQ_FILTER == 2
      const int col = i / 15, k = i - col * 15 - POLY_G0 * 5;
      if (k >= 0 && k < 5) {
        tile_f[k * PQ_TFS + col] = __double2float_rn(x);
        fm = fmaxf(fm, __double2float_ru(fabs(x)));
      }
#endif
    }
    // the columns of the last column step past the end of a ragged tile: +inf, they do not pass
#if PQ_FILTER == 2
    for (int i = ncol + threadIdx.x; i < ncol_pad; i += PC_THREADS)
#pragma unroll
      for (int k = 0; k < 5; ++k) tile_f[k * PQ_TFS + i] = INFINITY;
    fm = __uint_as_float(__reduce_max_sync(0xffffffffu, __float_as_uint(fm)));   // non-negative floats order like their bits
    if (lane == 0) atomicMax(&tile_fmax[it & 1], __float_as_uint(fm));
#else
    for (int i = ncol * 15 + threadIdx.x; i < ncol_pad * 15; i += PC_THREADS) tile[i] = INFINITY;
#endif
    __syncthreads();
#if PQ_FILTER == 2
    {
      const float f = fmaxf(f_row, __uint_as_float(tile_fmax[it & 1]));
      if (threadIdx.x == 0) tile_fmax[(it + 1) & 1] = 0u;   // the next tile's slot; its atomics come after the next barrier
      f_ok = f < 1e15f;                                        // f * f stays far from float overflow (and f is not NaN)
      margin = 81.0 * 5.9604644775390625e-8 * PQ_LOOSE * (double)f * (double)f;
      thr_filter = pq_float_threshold(thr, margin, f_ok);
    }
#endif
    for (int j0 = 0; j0 < ncol; j0 += PQ_UNROLL) {
#if PQ_FILTER == 2
      float g0[PQ_UNROLL];
#pragma unroll
      for (int u4 = 0; u4 < PQ_UNROLL; u4 += 4) {
        float2 s01, s23;
#pragma unroll
        for (int k = 0; k < 5; ++k) {
          const float4 bv = *reinterpret_cast<const float4*>(tile_f + k * PQ_TFS + j0 + u4);
          const float2 d01 = __fadd2_rn(make_float2(bv.x, bv.y), na[k]);
          const float2 d23 = __fadd2_rn(make_float2(bv.z, bv.w), na[k]);
          s01 = (k == 0) ? __fmul2_rn(d01, d01) : __ffma2_rn(d01, d01, s01);
          s23 = (k == 0) ? __fmul2_rn(d23, d23) : __ffma2_rn(d23, d23, s23);
        }
        g0[u4] = s01.x;
        g0[u4 + 1] = s01.y;
        g0[u4 + 2] = s23.x;
        g0[u4 + 3] = s23.y;
      }
#else
      double g0[PQ_UNROLL];
#pragma unroll
      for (int u = 0; u < PQ_UNROLL; ++u) {
        const double* b = tile + (j0 + u) * 15;
        double acc = 0.0;
#pragma unroll
        for (int k = 0; k < 5; ++k) {
          const double d = __dsub_rn(a0[k], b[POLY_G0 * 5 + k]);
          acc = (k == 0) ? __dmul_rn(d, d) : __fma_rn(d, d, acc);
        }
        g0[u] = acc;
      }
#endif
#pragma unroll
      for (int u = 0; u < PQ_UNROLL; ++u) {
        const bool pass = !(g0[u] >= thr_filter);
        const unsigned m = __ballot_sync(0xffffffffu, pass);
        if (pass) my_rc[q_n + __popc(m & ((1u << lane) - 1u))] = (lane << 16) | (j0 + u);
        q_n += __popc(m);
      }
      if (q_n >= 32) {   // one call site: dense batches from the front, the (< 32) rest moves up
        __syncwarp();
        int first = 0;
        while (q_n - first >= 32) {
          run_batch(first, 32, c0, ncol);
          first += 32;
        }
        const int rest = q_n - first;
        int mr = 0;
        if (lane < rest) mr = my_rc[first + lane];
        __syncwarp();
        if (lane < rest) my_rc[lane] = mr;
        q_n = rest;
        __syncwarp();
      }
    }
    (PQ_FILTER == 2 ? flops_f32 : flops) += live ? 14u * (unsigned)ncol : 0u;
    // the tile is about to be overwritten: drain what is left
    if (q_n > 0) {
      run_batch(0, q_n, c0, ncol);
      q_n = 0;
      __syncwarp();
    }
  }
  if (live) {
    argmin[pp.out_row + r] = best_c;
    mindist[pp.out_row + r] = best;
  }
  // flops executed by this warp (lane-level: pruned terms are not counted): [0] FP64, [1] single precision
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    flops += __shfl_xor_sync(0xffffffffu, flops, o);
    flops_f32 += __shfl_xor_sync(0xffffffffu, flops_f32, o);
  }
  if (lane == 0 && executed_flops) {
    atomicAdd(executed_flops, (unsigned long long)flops);
    if (flops_f32) atomicAdd(executed_flops + 1, (unsigned long long)flops_f32);
  }
}

// FP64 parity kernel, shift 0, with a SINGLE-PRECISION SCREEN in front (round 2, default).  What the
// reference keeps of a row is its first minimum; a column matters only if its distance is below the
// row's best so far, which happens O(log P2) times per row.  So every (row, column) cell is first
// evaluated in packed single precision -- all 15 properties, columns feature-major in shared memory,
// four columns per LDS.128, two columns per FADD2 / FFMA2, PS_ROWS rows per thread sharing the loads --
// against a threshold that makes the float sum a SAFE screen (below), and only the cells that pass are
// queued per warp (row slot, absolute column) and evaluated 32 at a time, from global memory, in FP64
// and in exactly the operations and the order of poly_dist2_f64<0>; improvements go to the lane that
// owns the row in ascending column order per row, which applies the reference's sequential rule.
// The result is the FP64 one bit for bit; the float arithmetic only decides what FP64 looks at.
//
// Safety of the screen.  u = 2^-24, F >= abs of every feature of the row and of the column tile,
// delta_k the real differences, e_k = fl32(fl32(b_k) - fl32(a_k)): abs(e_k - delta_k) <= 4.0000003 u F and
// abs(delta_k) <= 2 F give sum e_k^2 <= sum delta_k^2 + 240.0002 u F^2; the float accumulation (one FMUL,
// fourteen FFMA, terms >= 0) gives S_f <= (sum e_k^2)(1 + 15.0001 u).  A cell is dropped iff S_f >= T_f
// with T_f >= (thr (1 + 2^-48) + 241 u F^2 + 1e-36)(1 + 17 u), hence sum delta_k^2 >= thr (1 + 2^-48), and
// the FP64 value (nine roundings: >= (1 - 2^-49.8) sum delta_k^2) is >= thr: exactly a cell the
// reference's evaluation would not have taken.  Rounding towards a smaller S_f (underflow) only keeps
// more cells; the 1e-36 covers denormal rounding.  Where F^2 could leave the float range (or is NaN) T_f
// is NaN and every cell passes.  NaN sums pass too and are settled in FP64.
#ifndef PS_ROWS
#define PS_ROWS 2        // rows per thread (they share the column loads)
#endif
#ifndef PS_TILE_COLS
#define PS_TILE_COLS 128
#endif
#ifndef PS_UNROLL
#define PS_UNROLL 4      // columns per step
#endif
#ifndef PS_MIN_BLOCKS
#define PS_MIN_BLOCKS 7   // 72 registers (the FP64 batch is the rare path)
#endif
static_assert(PS_UNROLL % 4 == 0 && PS_TILE_COLS % PS_UNROLL == 0, "column steps are whole groups of four");
static_assert(PS_ROWS >= 1 && PS_ROWS <= 4, "row slot is packed into 8 bits");
constexpr int PS_TFS = PS_TILE_COLS + 4;             // floats per feature of the tile (16-byte aligned rows)
constexpr int PS_SLOTS = 32 * PS_ROWS;               // rows per warp
constexpr int PS_CTA_ROWS = PC_THREADS * PS_ROWS;    // rows per block
constexpr int PS_QUEUE = PS_SLOTS * PS_UNROLL + 32;  // 31 left over + everything one step can add
constexpr int PS_MAX_COLS = 1 << 24;                 // column index bits of a queue entry
constexpr int PS_AUTO_MIN_COLS = 8192;               // below this many columns the first-group queue kernel is the default

__device__ __forceinline__ float ps_float_threshold(double thr, double margin, bool f_ok) {
  const double t = (thr * (1.0 + 3.552713678800501e-15) + margin + 1e-36) * (1.0 + 17.0 * 5.9604644775390625e-8);
  return f_ok ? __double2float_ru(t) : __int_as_float(0x7fc00000);
}

__global__ void __launch_bounds__(PC_THREADS, PS_MIN_BLOCKS)
poly_compare_f64_screen_kernel(const PolyPair* __restrict__ pairs, int* __restrict__ argmin,
                               double* __restrict__ mindist, unsigned long long* __restrict__ executed_flops) {
  __shared__ __align__(16) float tile_f[15 * PS_TFS];            // the tile's columns in single precision, feature-major
  __shared__ unsigned tile_fmax[2];                               // max abs feature of the tile (float bits), by tile parity
  __shared__ unsigned thr_lo[PC_THREADS * PS_ROWS], thr_hi[PC_THREADS * PS_ROWS];   // FP64 thresholds by row slot
  __shared__ int q_rc[PC_THREADS / 32][PS_QUEUE];                 // row slot << 24 | column
  __shared__ int q_cnt[PC_THREADS / 32];
  const PolyPair pp = pairs[blockIdx.x];
  const int row0 = blockIdx.y * PS_CTA_ROWS;
  if (row0 >= pp.p1) return;
  constexpr unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int warp_row0 = row0 + wid * PS_SLOTS;    // row slot s = w * 32 + lane of this warp is row warp_row0 + s
  float na[PS_ROWS][15];                          // -fl32(feature)
  float f_row[PS_ROWS], thr_f[PS_ROWS];
  double thr[PS_ROWS], best[PS_ROWS], margin[PS_ROWS];
  int best_c[PS_ROWS];
  bool f_ok[PS_ROWS];
  unsigned* my_thr_lo = thr_lo + wid * PS_SLOTS;
  unsigned* my_thr_hi = thr_hi + wid * PS_SLOTS;
  int* my_rc = q_rc[wid];
#pragma unroll
  for (int w = 0; w < PS_ROWS; ++w) {
    const int r = warp_row0 + w * 32 + lane;
    const bool live = r < pp.p1;
    f_row[w] = 0.0f;
#pragma unroll
    for (int k = 0; k < 15; ++k) {
      const double x = live ? pp.f1[(long long)r * 15 + k] : 0.0;
      na[w][k] = -__double2float_rn(x);
      f_row[w] = fmaxf(f_row[w], __double2float_ru(fabs(x)));
    }
    thr[w] = live ? INFINITY : -1.0;             // a dead row: nothing it queues ever qualifies
    best[w] = INFINITY;
    best_c[w] = 0;
    thr_f[w] = 0.0f;                             // set per tile
    margin[w] = 0.0;
    f_ok[w] = true;
    my_thr_lo[w * 32 + lane] = (unsigned)__double2loint(thr[w]);
    my_thr_hi[w * 32 + lane] = (unsigned)__double2hiint(thr[w]);
  }
  if (threadIdx.x < 2) tile_fmax[threadIdx.x] = 0u;
  if (lane == 0) q_cnt[wid] = 0;
  unsigned flops = 0, flops_f32 = 0;

  // evaluates the queued entries [first, first + n), n <= 32, in FP64
  auto run_batch = [&](int first, int n) {
    double v = 0.0;
    int slot = 0, col = 0;
    bool cand = false;
    if (lane < n) {
      const int rc = my_rc[first + lane];
      slot = (int)((unsigned)rc >> 24);
      col = rc & (PS_MAX_COLS - 1);
    }
    const int row = warp_row0 + slot;
    if (lane < n && row < pp.p1 && col < pp.p2) {   // (dead rows and padding columns get here only through NaNs)
      double a[15];
#pragma unroll
      for (int k = 0; k < 15; ++k) a[k] = pp.f1[(long long)row * 15 + k];
      const double* b = pp.f2 + (long long)col * 15;
      const double t = __hiloint2double((int)my_thr_hi[slot], (int)my_thr_lo[slot]);
      v = poly_dist2_f64<0>(a, b);
      flops += 45;
      cand = v < t;
    }
    // improvements go to the lane that owns the row, in ascending entry (= column) order
    unsigned todo = __ballot_sync(FULL, cand);
    while (todo) {
      const int src = __ffs((int)todo) - 1;
      todo &= todo - 1;
      const int o_slot = __shfl_sync(FULL, slot, src);
      const int o_col = __shfl_sync(FULL, col, src);
      const double o_v = __shfl_sync(FULL, v, src);
      if (lane == (o_slot & 31)) {
#pragma unroll
        for (int w = 0; w < PS_ROWS; ++w)
          if ((o_slot >> 5) == w && o_v < thr[w]) {
            const double sv = sqrt(o_v);
            if (sv < best[w]) {
              best[w] = sv;
              best_c[w] = o_col;
            }
            thr[w] = o_v;
            thr_f[w] = ps_float_threshold(o_v, margin[w], f_ok[w]);
            my_thr_lo[o_slot] = (unsigned)__double2loint(o_v);
            my_thr_hi[o_slot] = (unsigned)__double2hiint(o_v);
          }
      }
      __syncwarp();
    }
  };

  for (int c0 = 0, it = 0; c0 < pp.p2; c0 += PS_TILE_COLS, ++it) {
    const int ncol = min(PS_TILE_COLS, pp.p2 - c0);
    const int ncol_pad = ((ncol + PS_UNROLL - 1) / PS_UNROLL) * PS_UNROLL;
    __syncthreads();
    float fm = 0.0f;
    for (int i = threadIdx.x; i < ncol * 15; i += PC_THREADS) {
      const double x = pp.f2[(long long)c0 * 15 + i];
      const int col = i / 15, k = i - col * 15;
      tile_f[k * PS_TFS + col] = __double2float_rn(x);
      fm = fmaxf(fm, __double2float_ru(fabs(x)));
    }
    // the columns of the last step past the end of a ragged tile: +inf, dropped by every finite threshold
    for (int i = ncol + threadIdx.x; i < ncol_pad; i += PC_THREADS)
#pragma unroll
      for (int k = 0; k < 15; ++k) tile_f[k * PS_TFS + i] = INFINITY;
    fm = __uint_as_float(__reduce_max_sync(FULL, __float_as_uint(fm)));   // non-negative floats order like their bits
    if (lane == 0) atomicMax(&tile_fmax[it & 1], __float_as_uint(fm));
    __syncthreads();
    {
      const float f_tile = __uint_as_float(tile_fmax[it & 1]);
      if (threadIdx.x == 0) tile_fmax[(it + 1) & 1] = 0u;   // the next tile's slot; its atomics come after the next barrier
#pragma unroll
      for (int w = 0; w < PS_ROWS; ++w) {
        const float f = fmaxf(f_row[w], f_tile);
        f_ok[w] = f < 1e15f;                                 // f * f stays far inside the float range (and f is not NaN)
        margin[w] = 241.0 * 5.9604644775390625e-8 * (1.0 + 9.094947017729282e-13) * (double)f * (double)f;
        thr_f[w] = ps_float_threshold(thr[w], margin[w], f_ok[w]);
      }
    }
    for (int j0 = 0; j0 < ncol; j0 += PS_UNROLL) {
      float2 s[PS_ROWS][PS_UNROLL / 2];
#pragma unroll
      for (int k = 0; k < 15; ++k) {
#pragma unroll
        for (int u4 = 0; u4 < PS_UNROLL / 4; ++u4) {
          const float4 bv = *reinterpret_cast<const float4*>(tile_f + k * PS_TFS + j0 + 4 * u4);
#pragma unroll
          for (int w = 0; w < PS_ROWS; ++w) {
            const float2 nk = make_float2(na[w][k], na[w][k]);
            const float2 d01 = __fadd2_rn(make_float2(bv.x, bv.y), nk);
            const float2 d23 = __fadd2_rn(make_float2(bv.z, bv.w), nk);
            s[w][2 * u4] = (k == 0) ? __fmul2_rn(d01, d01) : __ffma2_rn(d01, d01, s[w][2 * u4]);
            s[w][2 * u4 + 1] = (k == 0) ? __fmul2_rn(d23, d23) : __ffma2_rn(d23, d23, s[w][2 * u4 + 1]);
          }
        }
      }
      bool any = false;
#pragma unroll
      for (int w = 0; w < PS_ROWS; ++w)
#pragma unroll
        for (int h = 0; h < PS_UNROLL / 2; ++h)
          any = any || !(s[w][h].x >= thr_f[w]) || !(s[w][h].y >= thr_f[w]);
      if (__any_sync(FULL, any)) {
        // rare: queue the cells that passed, each row's in ascending column order
#pragma unroll
        for (int h = 0; h < PS_UNROLL / 2; ++h)
#pragma unroll
          for (int e = 0; e < 2; ++e)
#pragma unroll
            for (int w = 0; w < PS_ROWS; ++w) {
              const float sv = e ? s[w][h].y : s[w][h].x;
              if (!(sv >= thr_f[w]))
                my_rc[atomicAdd(&q_cnt[wid], 1)] = ((w * 32 + lane) << 24) | (c0 + j0 + 2 * h + e);
            }
        __syncwarp();
        const int n = *reinterpret_cast<volatile int*>(&q_cnt[wid]);
        if (n >= 32) {   // dense batches from the front, the (< 32) rest moves up
          int first = 0;
          while (n - first >= 32) {
            run_batch(first, 32);
            first += 32;
          }
          const int rest = n - first;
          int mr = 0;
          if (lane < rest) mr = my_rc[first + lane];
          __syncwarp();
          if (lane < rest) my_rc[lane] = mr;
          if (lane == 0) q_cnt[wid] = rest;
          __syncwarp();
        }
      }
    }
#pragma unroll
    for (int w = 0; w < PS_ROWS; ++w) flops_f32 += (warp_row0 + w * 32 + lane < pp.p1) ? 44u * (unsigned)ncol : 0u;
  }
  __syncwarp();
  {
    const int n = *reinterpret_cast<volatile int*>(&q_cnt[wid]);
    if (n > 0) run_batch(0, n);
  }
#pragma unroll
  for (int w = 0; w < PS_ROWS; ++w) {
    const int r = warp_row0 + w * 32 + lane;
    if (r < pp.p1) {
      argmin[pp.out_row + r] = best_c[w];
      mindist[pp.out_row + r] = best[w];
    }
  }
  // flops executed by this warp: [0] FP64, [1] single precision (44 per screened cell)
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    flops += __shfl_xor_sync(FULL, flops, o);
    flops_f32 += __shfl_xor_sync(FULL, flops_f32, o);
  }
  if (lane == 0 && executed_flops) {
    atomicAdd(executed_flops, (unsigned long long)flops);
    atomicAdd(executed_flops + 1, (unsigned long long)flops_f32);
  }
}

// FP32 fast mode (non-parity): features converted on the fly, FMA accumulation.
template <bool MIN_SHIFTS>
__global__ void __launch_bounds__(PC_THREADS)
poly_compare_f32_kernel(const PolyPair* __restrict__ pairs, int* __restrict__ argmin,
                        double* __restrict__ mindist) {
  __shared__ __align__(16) float tile[PC_TILE * 15];
  const PolyPair pp = pairs[blockIdx.x];
  const int row0 = blockIdx.y * PC_THREADS;
  if (row0 >= pp.p1) return;
  const int r = row0 + threadIdx.x;
  const bool live = r < pp.p1;
  float a[15];
#pragma unroll
  for (int k = 0; k < 15; ++k) a[k] = live ? (float)pp.f1[(long long)r * 15 + k] : 0.f;
  float best = INFINITY;
  int best_c = 0;
  for (int c0 = 0; c0 < pp.p2; c0 += PC_TILE) {
    const int ncol = min(PC_TILE, pp.p2 - c0);
    __syncthreads();
    for (int i = threadIdx.x; i < ncol * 15; i += PC_THREADS)
      tile[i] = (float)pp.f2[(long long)c0 * 15 + i];
    __syncthreads();
    if (live) {
      for (int j = 0; j < ncol; ++j) {
        const float* b = tile + j * 15;
        float v = poly_dist2_f32<0>(a, b);
        if (MIN_SHIFTS) {
          v = fminf(v, poly_dist2_f32<1>(a, b));
          v = fminf(v, poly_dist2_f32<2>(a, b));
          v = fminf(v, poly_dist2_f32<3>(a, b));
          v = fminf(v, poly_dist2_f32<4>(a, b));
        }
        if (v < best) {
          best = v;
          best_c = c0 + j;
        }
      }
    }
  }
  if (live) {
    argmin[pp.out_row + r] = best_c;
    mindist[pp.out_row + r] = (double)sqrtf(best);
  }
}

constexpr int PF_THREADS = 256;

__global__ void __launch_bounds__(PF_THREADS)
poly_count_kernel(const PolyPair* __restrict__ pairs, const double* __restrict__ mindist,
                  double threshold, int* __restrict__ counts) {
  const PolyPair pp = pairs[blockIdx.x];
  int local = 0;
  for (int r = threadIdx.x; r < pp.p1; r += PF_THREADS)
    local += (pp.p2 > 0 && mindist[pp.out_row + r] < threshold) ? 1 : 0;
  __shared__ int ws[PF_THREADS / 32];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) local += __shfl_xor_sync(0xffffffffu, local, o);
  if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = local;
  __syncthreads();
  if (threadIdx.x == 0) {
    int t = 0;
    for (int i = 0; i < PF_THREADS / 32; ++i) t += ws[i];
    counts[blockIdx.x] = t;
  }
}

__global__ void __launch_bounds__(PF_THREADS)
poly_compact_kernel(const PolyPair* __restrict__ pairs, const int* __restrict__ argmin,
                    const double* __restrict__ mindist, double threshold,
                    const long long* __restrict__ offsets, int* __restrict__ m1,
                    int* __restrict__ m2, double* __restrict__ md) {
  const PolyPair pp = pairs[blockIdx.x];
  __shared__ int warp_cnt[PF_THREADS / 32];
  __shared__ long long run;
  if (threadIdx.x == 0) run = offsets[blockIdx.x];
  __syncthreads();
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  for (int base = 0; base < pp.p1; base += PF_THREADS) {
    const int r = base + threadIdx.x;
    double d = 0.0;
    bool keep = false;
    if (r < pp.p1 && pp.p2 > 0) {
      d = mindist[pp.out_row + r];
      keep = d < threshold;
    }
    const unsigned ballot = __ballot_sync(0xffffffffu, keep);
    if (lane == 0) warp_cnt[wid] = __popc(ballot);
    __syncthreads();
    int before = 0, total = 0;
#pragma unroll
    for (int i = 0; i < PF_THREADS / 32; ++i) {
      const int c = warp_cnt[i];
      before += (i < wid) ? c : 0;
      total += c;
    }
    if (keep) {
      const long long pos = run + before + __popc(ballot & ((1u << lane) - 1u));
      m1[pos] = r;
      m2[pos] = argmin[pp.out_row + r];
      md[pos] = d;
    }
    __syncthreads();
    if (threadIdx.x == 0) run += total;
    __syncthreads();
  }
}

}  // namespace ssp

// =============================================================================================

#include "polygon_tensor.cuh"

// Launches the tensor-core screen and the exact part for the pairs of this call (polygon_tensor.cuh).
// Everything is queued on ctx->stream; flags and statistics are read by the caller behind its one
// synchronisation.  `counted`: the device counter holds the FP64 flops of the exact part.
static cudaError_t launch_poly_compare_tensor(ssp_ctx* ctx, const std::vector<ssp::PolyPair>& pp,
                                              const int32_t* pairs, int n_pairs, long long rows) {
  using namespace ssp;
  cudaError_t e;
  // operand banks of the images this call touches
  std::vector<int> slot_of(ctx->poly.size(), -1);
  std::vector<PtBankImg> imgs;
  std::vector<long long> off, row0;
  long long bytes = 0, n_rows = 0;
  int max_pad = 0;
  for (int p = 0; p < n_pairs; ++p)
    for (int s = 0; s < 2; ++s) {
      const int i = pairs[2 * p + s];
      if (slot_of[(size_t)i] >= 0) continue;
      const long long count = ctx->poly[(size_t)i].count;
      PtBankImg im;
      im.feat = ctx->poly[(size_t)i].feat.as<double>();
      im.count = (int)count;
      im.rows_pad = (int)((count + PT_PAD - 1) / PT_PAD * PT_PAD);
      im.out_off = bytes;
      im.row0 = n_rows;
      im.slot = (int)imgs.size();
      im.pad_ = 0;
      slot_of[(size_t)i] = im.slot;
      bytes += 2ll * im.rows_pad * PT_ROW_BYTES;
      n_rows += count;
      max_pad = im.rows_pad > max_pad ? im.rows_pad : max_pad;
      imgs.push_back(im);
    }
  if ((e = ctx->t_bank.ensure((size_t)bytes + 1024)) != cudaSuccess) return e;
  if ((e = ctx->t_na2.ensure((size_t)(n_rows + 1) * 8)) != cudaSuccess) return e;
  if ((e = ctx->t_f.ensure(imgs.size() * 4 + 16)) != cudaSuccess) return e;   // + statistics word behind the table
  if ((e = ctx->t_imgs.ensure(imgs.size() * sizeof(PtBankImg))) != cudaSuccess) return e;
  if ((e = ctx->t_flags.ensure(2 * sizeof(int))) != cudaSuccess) return e;   // flags, number of kept candidates
  if ((e = cudaMemsetAsync(ctx->t_f.p, 0, imgs.size() * 4 + 16, ctx->stream)) != cudaSuccess) return e;
  if ((e = cudaMemsetAsync(ctx->t_flags.p, 0, 2 * sizeof(int), ctx->stream)) != cudaSuccess) return e;
  // pageable sources: the runtime stages them before returning
  if ((e = cudaMemcpyAsync(ctx->t_imgs.p, imgs.data(), imgs.size() * sizeof(PtBankImg), cudaMemcpyHostToDevice,
                           ctx->stream)) != cudaSuccess)
    return e;
  // work list, one entry per pair (pairs without rows or columns own no tile)
  std::vector<PtWork> tw((size_t)n_pairs);
  std::vector<const double*> na2ptr((size_t)n_pairs);
  int total = 0;
  for (int p = 0; p < n_pairs; ++p) {
    const PtBankImg& a = imgs[(size_t)slot_of[(size_t)pairs[2 * p]]];
    const PtBankImg& b = imgs[(size_t)slot_of[(size_t)pairs[2 * p + 1]]];
    PtWork& w = tw[(size_t)p];
    w.a_exp = ctx->t_bank.as<uint8_t>() + a.out_off;
    w.b_exp = ctx->t_bank.as<uint8_t>() + b.out_off + (size_t)b.rows_pad * PT_ROW_BYTES;
    w.f_a = ctx->t_f.as<float>() + a.slot;
    w.f_b = ctx->t_f.as<float>() + b.slot;
    w.p1 = pp[(size_t)p].p1;
    w.p2 = pp[(size_t)p].p2;
    w.a_rows_pad = a.rows_pad;
    w.b_rows_pad = b.rows_pad;
    w.tile_begin = total;
    w.pair = p;
    w.out_row = pp[(size_t)p].out_row;
    na2ptr[(size_t)p] = ctx->t_na2.as<double>() + a.row0;
    if (w.p1 > 0 && w.p2 > 0) total += (w.p1 + PT_M - 1) / PT_M;
  }
  if ((e = ctx->t_work.ensure(tw.size() * sizeof(PtWork))) != cudaSuccess) return e;
  if ((e = ctx->t_na2ptr.ensure(na2ptr.size() * sizeof(double*))) != cudaSuccess) return e;
  if ((e = cudaMemcpyAsync(ctx->t_work.p, tw.data(), tw.size() * sizeof(PtWork), cudaMemcpyHostToDevice,
                           ctx->stream)) != cudaSuccess)
    return e;
  if ((e = cudaMemcpyAsync(ctx->t_na2ptr.p, na2ptr.data(), na2ptr.size() * sizeof(double*), cudaMemcpyHostToDevice,
                           ctx->stream)) != cudaSuccess)
    return e;
  const int grid = total < ctx->sm_count ? (total > 0 ? total : 1) : ctx->sm_count;
  // candidate pool: ~ln(P2) + a few per row are expected; 24 per row and a floor per CTA, split evenly
  // (pool positions are 32-bit: a call beyond that simply overflows the pool and goes to the CUDA-core kernels)
  const unsigned long long pool_want = 24ull * (unsigned long long)rows + 8192ull * (unsigned)grid;
  const unsigned long long pool_total = pool_want < 0xF0000000ull ? pool_want : 0xF0000000ull;
  const unsigned seg_cap = (unsigned)(pool_total / (unsigned)grid);
  if ((e = ctx->t_pool.ensure((size_t)seg_cap * grid * sizeof(PtCand))) != cudaSuccess) return e;
  if ((e = ctx->t_sv.ensure((size_t)seg_cap * grid * 8)) != cudaSuccess) return e;
  if ((e = ctx->t_keep.ensure((size_t)seg_cap * grid * 4)) != cudaSuccess) return e;   // worst case: every candidate kept
  if ((e = ctx->t_counts.ensure((size_t)grid * 4)) != cudaSuccess) return e;
  if ((e = ctx->t_rowkey.ensure((size_t)(rows + 1) * 12)) != cudaSuccess) return e;   // row keys, then the rows' final thetas
  unsigned* row_theta = reinterpret_cast<unsigned*>(ctx->t_rowkey.as<unsigned long long>() + (rows + 1));
  if ((e = cudaMemsetAsync(ctx->t_counts.p, 0, (size_t)grid * 4, ctx->stream)) != cudaSuccess) return e;

  if ((e = cudaEventRecord(ctx->ev0, ctx->stream)) != cudaSuccess) return e;
  {
    const long long threads = (long long)max_pad * 16;
    poly_tf32_bank_kernel<<<dim3((unsigned)((threads + 255) / 256), (unsigned)imgs.size()), 256, 0, ctx->stream>>>(
        ctx->t_imgs.as<PtBankImg>(), ctx->t_bank.as<uint8_t>(), ctx->t_na2.as<double>(), ctx->t_f.as<unsigned>());
    if ((e = cudaGetLastError()) != cudaSuccess) return e;
  }
  poly_cand_init_kernel<<<(unsigned)((rows + 255) / 256), 256, 0, ctx->stream>>>(
      ctx->t_rowkey.as<unsigned long long>(), ctx->p_argmin.as<int>(), row_theta, rows);
  if ((e = cudaGetLastError()) != cudaSuccess) return e;
  if (total > 0) {
    if ((e = cudaFuncSetAttribute(poly_screen_tcgen05_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  (int)PT_SMEM_BYTES)) != cudaSuccess)
      return e;
    poly_screen_tcgen05_kernel<<<grid, PT_THREADS, PT_SMEM_BYTES, ctx->stream>>>(
        ctx->t_work.as<PtWork>(), n_pairs, total, ctx->t_pool.as<PtCand>(), seg_cap, ctx->t_counts.as<unsigned>(),
        row_theta, ctx->t_flags.as<int>());
    if ((e = cudaGetLastError()) != cudaSuccess) return e;
    unsigned* stats = ctx->t_f.as<unsigned>() + imgs.size();
    unsigned* n_keep = ctx->t_flags.as<unsigned>() + 1;
    poly_cand_filter_kernel<<<dim3(64, (unsigned)grid), 256, 0, ctx->stream>>>(
        ctx->p_pairs.as<PolyPair>(), ctx->t_pool.as<PtCand>(), seg_cap, ctx->t_counts.as<unsigned>(), row_theta,
        ctx->t_keep.as<unsigned>(), n_keep);
    if ((e = cudaGetLastError()) != cudaSuccess) return e;
    const unsigned dense_grid = (unsigned)ctx->sm_count * 8u;
    poly_cand_eval_kernel<<<dense_grid, 256, 0, ctx->stream>>>(
        ctx->p_pairs.as<PolyPair>(), ctx->t_work.as<PtWork>(), ctx->t_pool.as<PtCand>(), ctx->t_keep.as<unsigned>(),
        n_keep, ctx->t_na2ptr.as<const double*>(), ctx->t_sv.as<double>(), ctx->t_rowkey.as<unsigned long long>(),
        stats, ctx->t_flags.as<int>(), ctx->p_flops.as<unsigned long long>());
    if ((e = cudaGetLastError()) != cudaSuccess) return e;
    poly_cand_argmin_kernel<<<dense_grid, 256, 0, ctx->stream>>>(
        ctx->p_pairs.as<PolyPair>(), ctx->t_pool.as<PtCand>(), ctx->t_keep.as<unsigned>(), n_keep,
        ctx->t_sv.as<double>(), ctx->t_rowkey.as<unsigned long long>(), ctx->p_argmin.as<int>());
    if ((e = cudaGetLastError()) != cudaSuccess) return e;
  }
  poly_cand_finish_kernel<<<(unsigned)((rows + 255) / 256), 256, 0, ctx->stream>>>(
      ctx->t_rowkey.as<unsigned long long>(), ctx->p_argmin.as<int>(), ctx->p_mindist.as<double>(), rows);
  if ((e = cudaGetLastError()) != cudaSuccess) return e;
  if ((e = cudaEventRecord(ctx->ev1, ctx->stream)) != cudaSuccess) return e;
  ctx->p_launches += total > 0 ? 7 : 3;
  ctx->t_stats_off = (long long)imgs.size();
  ctx->t_grid = grid;
  return cudaSuccess;
}


extern "C" {

int ssp_upload_star_bank(ssp_ctx* ctx, const double* xy, const double* brightness,
                         const int64_t* star_offsets, int n_images, int img_w, int img_h,
                         double focal_px, int mem_kind) {
  using namespace ssp;
  if (!ctx) return SSP_ERR_INVALID;
  if (int rc = use_device(ctx)) return rc;
  SSP_REQUIRE(ctx, star_offsets != nullptr && n_images >= 0 && star_offsets[0] == 0,
              SSP_ERR_INVALID, "ssp_upload_star_bank: bad offsets");
  for (int i = 0; i < n_images; ++i)
    SSP_REQUIRE(ctx, star_offsets[i + 1] >= star_offsets[i] &&
                         star_offsets[i + 1] - star_offsets[i] < (1LL << 31) - 1,
                SSP_ERR_INVALID, "ssp_upload_star_bank: bad offsets at image %d", i);
  SSP_REQUIRE(ctx, img_w > 0 && img_h > 0 && focal_px > 0.0, SSP_ERR_INVALID,
              "ssp_upload_star_bank: image size and focal length must be positive");
  const long long n = star_offsets[n_images];
  SSP_REQUIRE(ctx, n == 0 || (xy != nullptr && brightness != nullptr), SSP_ERR_INVALID,
              "ssp_upload_star_bank: NULL star arrays");
  SSP_CUDA_TRY(ctx, ctx->s_xy.ensure((size_t)n * 16 + 16));
  SSP_CUDA_TRY(ctx, ctx->s_bright.ensure((size_t)n * 8 + 8));
  SSP_CUDA_TRY(ctx, ctx->s_rays.ensure((size_t)n * 24 + 24));
  if (n > 0) {
    const cudaMemcpyKind kind = to_device_kind(mem_kind);
    SSP_CUDA_TRY(ctx, cudaMemcpyAsync(ctx->s_xy.p, xy, (size_t)n * 16, kind, ctx->stream));
    SSP_CUDA_TRY(ctx, cudaMemcpyAsync(ctx->s_bright.p, brightness, (size_t)n * 8, kind, ctx->stream));
    // half sizes exactly as the reference: int(img_shape[1] / 2), int(img_shape[0] / 2)
    const double half_w = (double)(img_w / 2), half_h = (double)(img_h / 2);
    star_rays_kernel<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(
        ctx->s_xy.as<double>(), n, half_w, half_h, focal_px, ctx->s_rays.as<double>());
    SSP_CUDA_TRY(ctx, cudaGetLastError());
  }
  ctx->star_off.assign(star_offsets, star_offsets + n_images + 1);
  ctx->img_w = img_w;
  ctx->img_h = img_h;
  ctx->focal = focal_px;
  // polygon banks measured earlier stay valid (they hold their own copies); re-measure an
  // image after changing its stars.
  if (ctx->poly.size() < (size_t)n_images) ctx->poly.resize((size_t)n_images);
  return SSP_OK;
}

int ssp_fetch_star_rays(ssp_ctx* ctx, double* rays, int64_t capacity_stars, int mem_kind) {
  using namespace ssp;
  if (!ctx) return SSP_ERR_INVALID;
  if (int rc = use_device(ctx)) return rc;
  SSP_REQUIRE(ctx, !ctx->star_off.empty(), SSP_ERR_STATE, "ssp_fetch_star_rays: no star bank");
  const long long n = ctx->star_off.back();
  SSP_REQUIRE(ctx, capacity_stars >= n, SSP_ERR_CAPACITY, "ssp_fetch_star_rays: capacity too small");
  SSP_REQUIRE(ctx, rays != nullptr || n == 0, SSP_ERR_INVALID, "ssp_fetch_star_rays: NULL buffer");
  if (n > 0)
    SSP_CUDA_TRY(ctx, cudaMemcpyAsync(rays, ctx->s_rays.p, (size_t)n * 24, from_device_kind(mem_kind),
                                      ctx->stream));
  SSP_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  return SSP_OK;
}

static int ensure_poly_image(ssp_ctx* ctx, int image, const char* who, bool need_stars) {
  // polygon banks can also be loaded without a star bank (ssp_polygon_upload_features)
  SSP_REQUIRE(ctx, image >= 0 && image < (1 << 20), SSP_ERR_INVALID, "%s: bad image index %d", who,
              image);
  if (need_stars)
    SSP_REQUIRE(ctx, image < (int)ctx->star_off.size() - 1, SSP_ERR_INVALID,
                "%s: image %d is not in the star bank", who, image);
  if ((size_t)image >= ctx->poly.size()) ctx->poly.resize((size_t)image + 1);
  return SSP_OK;
}

// Completes every ssp_polygon_measure call issued with n_valid == NULL: one stream synchronise,
// then counts and domain flags are read from their pinned slots.
static int finish_pending_measures(ssp_ctx* ctx, const char* who) {
  using namespace ssp;
  if (!ctx->p_any_pending) return SSP_OK;
  SSP_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  ctx->p_any_pending = false;
  int bad = -1, n_bad = 0;
  for (size_t i = 0; i < ctx->poly.size(); ++i) {
    PolyImage& pi = ctx->poly[i];
    if (!pi.pending) continue;
    pi.pending = false;
    pi.count = ctx->p_slots[2 * i];
    if (ctx->p_slots[2 * i + 1] > 0 && bad < 0) {
      bad = (int)i;
      n_bad = (int)ctx->p_slots[2 * i + 1];
    }
  }
  if (bad >= 0)
    return fail(ctx, SSP_ERR_DOMAIN,
                "%s: %d candidate(s) of image %d gave |acos argument| > 1 (the reference raises "
                "ValueError at custom_bf_polygon_matcher.py:134)", who, n_bad, bad);
  return SSP_OK;
}

int ssp_polygon_measure(ssp_ctx* ctx, int image, const int32_t* candidates, int64_t n_candidates,
                        double min_side_px, int mem_kind, uint8_t* status, int64_t* n_valid) {
  using namespace ssp;
  if (!ctx) return SSP_ERR_INVALID;
  if (int rc = use_device(ctx)) return rc;
  SSP_REQUIRE(ctx, !ctx->star_off.empty(), SSP_ERR_STATE, "ssp_polygon_measure: no star bank");
  if (int rc = ensure_poly_image(ctx, image, "ssp_polygon_measure", true)) return rc;
  SSP_REQUIRE(ctx, n_candidates >= 0 && n_candidates < (1LL << 31) - 1024, SSP_ERR_INVALID,
              "ssp_polygon_measure: bad candidate count");
  SSP_REQUIRE(ctx, n_candidates == 0 || candidates != nullptr, SSP_ERR_INVALID,
              "ssp_polygon_measure: NULL candidates");
  PolyImage& pi = ctx->poly[(size_t)image];
  const long long n = n_candidates;
  ctx->p_n_matches = -1;
  if (n == 0) {
    pi.count = 0;
    pi.pending = false;
    pi.has_raw = true;
    if (n_valid) *n_valid = 0;
    return SSP_OK;
  }
  const long long s0 = ctx->star_off[(size_t)image];
  StarView sv;
  sv.xy = ctx->s_xy.as<double>() + 2 * s0;
  sv.bright = ctx->s_bright.as<double>() + s0;
  sv.rays = ctx->s_rays.as<double>() + 3 * s0;
  sv.n = (int)(ctx->star_off[(size_t)image + 1] - s0);

  SSP_CUDA_TRY(ctx, ctx->p_cand.ensure((size_t)n * 20));
  SSP_CUDA_TRY(ctx, ctx->p_status.ensure((size_t)n + 16));
  SSP_CUDA_TRY(ctx, ctx->p_scan.ensure((size_t)n * 4 + (size_t)(n + 1) * 8 + 64));
  SSP_CUDA_TRY(ctx, ctx->p_tmp_idx.ensure((size_t)n * 20));
  SSP_CUDA_TRY(ctx, ctx->p_tmp_angle.ensure((size_t)n * 40));
  SSP_CUDA_TRY(ctx, ctx->p_tmp_side.ensure((size_t)n * 40));
  SSP_CUDA_TRY(ctx, ctx->p_tmp_bright.ensure((size_t)n * 40));
  SSP_CUDA_TRY(ctx, ctx->p_tmp_feat.ensure((size_t)n * 120));
  SSP_CUDA_TRY(ctx, pi.star_idx.ensure((size_t)n * 20));
  SSP_CUDA_TRY(ctx, pi.angle.ensure((size_t)n * 40));
  SSP_CUDA_TRY(ctx, pi.side.ensure((size_t)n * 40));
  SSP_CUDA_TRY(ctx, pi.bright.ensure((size_t)n * 40));
  SSP_CUDA_TRY(ctx, pi.feat.ensure((size_t)n * 120));
  // p_scan layout: [n_domain int][pad][offsets (n+1) int64][flags n int]
  int* d_ndomain = ctx->p_scan.as<int>();
  long long* d_pos = reinterpret_cast<long long*>(ctx->p_scan.as<char>() + 16);
  int* d_flags = reinterpret_cast<int*>(ctx->p_scan.as<char>() + 16 + (size_t)(n + 1) * 8);
  SSP_CUDA_TRY(ctx, cudaMemsetAsync(d_ndomain, 0, 16, ctx->stream));
  SSP_CUDA_TRY(ctx, cudaMemcpyAsync(ctx->p_cand.p, candidates, (size_t)n * 20,
                                    to_device_kind(mem_kind), ctx->stream));
  const unsigned blocks = (unsigned)((n + 127) / 128);
  poly_measure_kernel<<<blocks, 128, 0, ctx->stream>>>(
      sv, ctx->p_cand.as<int>(), n, min_side_px, ctx->p_status.as<unsigned char>(), d_flags,
      d_ndomain, ctx->p_tmp_idx.as<int>(), ctx->p_tmp_angle.as<double>(),
      ctx->p_tmp_side.as<double>(), ctx->p_tmp_bright.as<double>(), ctx->p_tmp_feat.as<double>());
  SSP_CUDA_TRY(ctx, cudaGetLastError());
  poly_scatter_kernel<<<blocks, 128, 0, ctx->stream>>>(
      d_flags, d_pos + n, n, ctx->p_tmp_idx.as<int>(), ctx->p_tmp_angle.as<double>(),
      ctx->p_tmp_side.as<double>(), ctx->p_tmp_bright.as<double>(), ctx->p_tmp_feat.as<double>(),
      pi.star_idx.as<int>(), pi.angle.as<double>(), pi.side.as<double>(), pi.bright.as<double>(),
      pi.feat.as<double>());
  SSP_CUDA_TRY(ctx, cudaGetLastError());
  constexpr int SLOT_IMAGES = 4096;
  if (!n_valid && !status && image < SLOT_IMAGES) {
    // no-wait form: the caller does not ask for the count now, so nothing here blocks; the
    // count and the domain flag travel to a pinned slot and are read by the next call that
    // needs them (ssp_polygon_count / _fetch / _match_pairs).  Scratch buffers are reused by
    // the next image in stream order.
    if (!ctx->p_slots) {
      SSP_CUDA_TRY(ctx, cudaHostAlloc(reinterpret_cast<void**>(&ctx->p_slots),
                                      (size_t)SLOT_IMAGES * 16, cudaHostAllocPortable));
      ctx->p_slot_capacity = SLOT_IMAGES;
    }
    ctx->p_slots[2 * image + 1] = 0;
    SSP_CUDA_TRY(ctx, cudaMemcpyAsync(&ctx->p_slots[2 * image], d_pos + n, 8, cudaMemcpyDeviceToHost,
                                      ctx->stream));
    SSP_CUDA_TRY(ctx, cudaMemcpyAsync(&ctx->p_slots[2 * image + 1], d_ndomain, 4, cudaMemcpyDeviceToHost,
                                      ctx->stream));
    pi.count = 0;
    pi.pending = true;
    pi.has_raw = true;
    ctx->p_any_pending = true;
    return SSP_OK;
  }
  pi.pending = false;
  long long count = 0;
  int n_domain = 0;
  SSP_CUDA_TRY(ctx, cudaMemcpyAsync(&count, d_pos + n, 8, cudaMemcpyDeviceToHost, ctx->stream));
  SSP_CUDA_TRY(ctx, cudaMemcpyAsync(&n_domain, d_ndomain, 4, cudaMemcpyDeviceToHost, ctx->stream));
  if (status)
    SSP_CUDA_TRY(ctx, cudaMemcpyAsync(status, ctx->p_status.p, (size_t)n, cudaMemcpyDeviceToHost,
                                      ctx->stream));
  SSP_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  pi.count = count;
  pi.has_raw = true;
  if (n_valid) *n_valid = count;
  if (n_domain > 0)
    return fail(ctx, SSP_ERR_DOMAIN,
                "ssp_polygon_measure: %d candidate(s) of image %d gave |acos argument| > 1 "
                "(the reference raises ValueError at custom_bf_polygon_matcher.py:134)",
                n_domain, image);
  return SSP_OK;
}

int ssp_polygon_measure_batch(ssp_ctx* ctx, const int32_t* images, int n_images, const int32_t* candidates,
                              const int64_t* cand_offsets, double min_side_px, int mem_kind) {
  using namespace ssp;
  if (!ctx) return SSP_ERR_INVALID;
  if (int rc = use_device(ctx)) return rc;
  SSP_REQUIRE(ctx, !ctx->star_off.empty(), SSP_ERR_STATE, "ssp_polygon_measure_batch: no star bank");
  SSP_REQUIRE(ctx, n_images >= 0 && (n_images == 0 || (images != nullptr && cand_offsets != nullptr)) &&
                       (n_images == 0 || cand_offsets[0] == 0),
              SSP_ERR_INVALID, "ssp_polygon_measure_batch: bad arguments");
  if (n_images == 0) return SSP_OK;
  constexpr int SLOT_IMAGES = 4096;
  const long long n = cand_offsets[n_images];
  SSP_REQUIRE(ctx, n >= 0 && n < (1LL << 31) - 1024 && (n == 0 || candidates != nullptr), SSP_ERR_INVALID,
              "ssp_polygon_measure_batch: bad candidate table");
  std::vector<MeasureImage> tab;
  int blocks = 0;
  for (int k = 0; k < n_images; ++k) {
    const int image = images[k];
    if (int rc = ensure_poly_image(ctx, image, "ssp_polygon_measure_batch", true)) return rc;
    SSP_REQUIRE(ctx, image < SLOT_IMAGES, SSP_ERR_INVALID, "ssp_polygon_measure_batch: image id %d >= %d", image, SLOT_IMAGES);
    for (int j = 0; j < k; ++j)
      SSP_REQUIRE(ctx, images[j] != image, SSP_ERR_INVALID, "ssp_polygon_measure_batch: image %d listed twice", image);
    const long long ni = cand_offsets[k + 1] - cand_offsets[k];
    SSP_REQUIRE(ctx, ni >= 0, SSP_ERR_INVALID, "ssp_polygon_measure_batch: offsets must be non-decreasing");
  }
  if (!ctx->p_slots) {
    SSP_CUDA_TRY(ctx, cudaHostAlloc(reinterpret_cast<void**>(&ctx->p_slots), (size_t)SLOT_IMAGES * 16,
                                    cudaHostAllocPortable));
    ctx->p_slot_capacity = SLOT_IMAGES;
  }
  ctx->p_n_matches = -1;
  NvtxRange range("kernel 2 (polygon measure, batch)");
  // scratch: as ssp_polygon_measure, sized for all images together
  const long long ns = n > 0 ? n : 1;
  SSP_CUDA_TRY(ctx, ctx->p_cand.ensure((size_t)ns * 20));
  SSP_CUDA_TRY(ctx, ctx->p_status.ensure((size_t)ns + 16));
  const size_t head = ((size_t)n_images * (16 + sizeof(MeasureImage)) + 63) / 64 * 64;
  SSP_CUDA_TRY(ctx, ctx->p_scan.ensure(head + (size_t)ns * 4 + 64));
  SSP_CUDA_TRY(ctx, ctx->p_tmp_idx.ensure((size_t)ns * 20));
  SSP_CUDA_TRY(ctx, ctx->p_tmp_angle.ensure((size_t)ns * 40));
  SSP_CUDA_TRY(ctx, ctx->p_tmp_side.ensure((size_t)ns * 40));
  SSP_CUDA_TRY(ctx, ctx->p_tmp_bright.ensure((size_t)ns * 40));
  SSP_CUDA_TRY(ctx, ctx->p_tmp_feat.ensure((size_t)ns * 120));
  // p_scan layout here: [slots n_images x 2 int64][table n_images x MeasureImage][flags n int]
  long long* d_slots = ctx->p_scan.as<long long>();
  MeasureImage* d_tab = reinterpret_cast<MeasureImage*>(ctx->p_scan.as<char>() + (size_t)n_images * 16);
  int* d_flags = reinterpret_cast<int*>(ctx->p_scan.as<char>() + head);
  for (int k = 0; k < n_images; ++k) {
    const int image = images[k];
    PolyImage& pi = ctx->poly[(size_t)image];
    const long long ni = cand_offsets[k + 1] - cand_offsets[k];
    const long long cap = ni > 0 ? ni : 1;
    SSP_CUDA_TRY(ctx, pi.star_idx.ensure((size_t)cap * 20));
    SSP_CUDA_TRY(ctx, pi.angle.ensure((size_t)cap * 40));
    SSP_CUDA_TRY(ctx, pi.side.ensure((size_t)cap * 40));
    SSP_CUDA_TRY(ctx, pi.bright.ensure((size_t)cap * 40));
    SSP_CUDA_TRY(ctx, pi.feat.ensure((size_t)cap * 120));
    const long long s0 = ctx->star_off[(size_t)image];
    MeasureImage m;
    m.sv.xy = ctx->s_xy.as<double>() + 2 * s0;
    m.sv.bright = ctx->s_bright.as<double>() + s0;
    m.sv.rays = ctx->s_rays.as<double>() + 3 * s0;
    m.sv.n = (int)(ctx->star_off[(size_t)image + 1] - s0);
    m.cand_off = cand_offsets[k];
    m.n = (int)ni;
    m.block_off = blocks;
    m.o_idx = pi.star_idx.as<int>();
    m.o_angle = pi.angle.as<double>();
    m.o_side = pi.side.as<double>();
    m.o_bright = pi.bright.as<double>();
    m.o_feat = pi.feat.as<double>();
    m.slot = d_slots + 2 * k;
    blocks += (int)((ni + 127) / 128);
    tab.push_back(m);
  }
  SSP_CUDA_TRY(ctx, cudaMemsetAsync(d_slots, 0, (size_t)n_images * 16, ctx->stream));
  SSP_CUDA_TRY(ctx, cudaMemcpyAsync(d_tab, tab.data(), tab.size() * sizeof(MeasureImage), cudaMemcpyHostToDevice,
                                    ctx->stream));   // pageable source: staged before the call returns
  if (n > 0) {
    SSP_CUDA_TRY(ctx, cudaMemcpyAsync(ctx->p_cand.p, candidates, (size_t)n * 20, to_device_kind(mem_kind), ctx->stream));
    poly_measure_batch_kernel<<<(unsigned)blocks, 128, 0, ctx->stream>>>(
        d_tab, n_images, ctx->p_cand.as<int>(), min_side_px, ctx->p_status.as<unsigned char>(), d_flags,
        ctx->p_tmp_idx.as<int>(), ctx->p_tmp_angle.as<double>(), ctx->p_tmp_side.as<double>(),
        ctx->p_tmp_bright.as<double>(), ctx->p_tmp_feat.as<double>());
    SSP_CUDA_TRY(ctx, cudaGetLastError());
    poly_scatter_batch_kernel<<<(unsigned)blocks, 128, 0, ctx->stream>>>(
        d_tab, n_images, d_flags, ctx->p_tmp_idx.as<int>(), ctx->p_tmp_angle.as<double>(),
        ctx->p_tmp_side.as<double>(), ctx->p_tmp_bright.as<double>(), ctx->p_tmp_feat.as<double>());
    SSP_CUDA_TRY(ctx, cudaGetLastError());
  }
  // counts and domain flags travel to the pinned slots; the next consumer reads them (no wait here)
  for (int k = 0; k < n_images; ++k) {
    const int image = images[k];
    PolyImage& pi = ctx->poly[(size_t)image];
    SSP_CUDA_TRY(ctx, cudaMemcpyAsync(&ctx->p_slots[2 * image], d_slots + 2 * k, 16, cudaMemcpyDeviceToHost, ctx->stream));
    pi.count = 0;
    pi.pending = true;
    pi.has_raw = true;
  }
  ctx->p_any_pending = true;
  return SSP_OK;
}

int ssp_polygon_count(ssp_ctx* ctx, int image, int64_t* n_polygons) {
  using namespace ssp;
  if (!ctx) return SSP_ERR_INVALID;
  if (int rc = use_device(ctx)) return rc;
  if (int rc = finish_pending_measures(ctx, "ssp_polygon_count")) return rc;
  SSP_REQUIRE(ctx, image >= 0 && (size_t)image < ctx->poly.size() && ctx->poly[(size_t)image].count >= 0,
              SSP_ERR_STATE, "ssp_polygon_count: image %d has no polygon bank", image);
  if (n_polygons) *n_polygons = ctx->poly[(size_t)image].count;
  return SSP_OK;
}

int ssp_polygon_fetch(ssp_ctx* ctx, int image, int32_t* star_idx, double* angle, double* side,
                      double* brightness, double* feat, int64_t capacity, int mem_kind) {
  using namespace ssp;
  if (!ctx) return SSP_ERR_INVALID;
  if (int rc = use_device(ctx)) return rc;
  if (int rc = finish_pending_measures(ctx, "ssp_polygon_fetch")) return rc;
  SSP_REQUIRE(ctx, image >= 0 && (size_t)image < ctx->poly.size() && ctx->poly[(size_t)image].count >= 0,
              SSP_ERR_STATE, "ssp_polygon_fetch: image %d has no polygon bank", image);
  PolyImage& pi = ctx->poly[(size_t)image];
  const long long n = pi.count;
  SSP_REQUIRE(ctx, capacity >= n, SSP_ERR_CAPACITY, "ssp_polygon_fetch: capacity %lld < %lld",
              (long long)capacity, n);
  SSP_REQUIRE(ctx, pi.has_raw || (angle == nullptr && side == nullptr && brightness == nullptr),
              SSP_ERR_STATE,
              "ssp_polygon_fetch: image %d was loaded with ssp_polygon_upload_features, raw "
              "measures are not available", image);
  if (n > 0) {
    const cudaMemcpyKind kind = from_device_kind(mem_kind);
    if (star_idx && pi.star_idx.p)
      SSP_CUDA_TRY(ctx, cudaMemcpyAsync(star_idx, pi.star_idx.p, (size_t)n * 20, kind, ctx->stream));
    if (angle) SSP_CUDA_TRY(ctx, cudaMemcpyAsync(angle, pi.angle.p, (size_t)n * 40, kind, ctx->stream));
    if (side) SSP_CUDA_TRY(ctx, cudaMemcpyAsync(side, pi.side.p, (size_t)n * 40, kind, ctx->stream));
    if (brightness)
      SSP_CUDA_TRY(ctx, cudaMemcpyAsync(brightness, pi.bright.p, (size_t)n * 40, kind, ctx->stream));
    if (feat) SSP_CUDA_TRY(ctx, cudaMemcpyAsync(feat, pi.feat.p, (size_t)n * 120, kind, ctx->stream));
  }
  // device destinations are stream-ordered with whatever the caller enqueues next on this stream
  if (mem_kind != SSP_MEM_DEVICE) SSP_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  return SSP_OK;
}

int ssp_polygon_upload_features(ssp_ctx* ctx, int image, const double* feat,
                                const int32_t* star_idx, int64_t n_polygons, int mem_kind) {
  using namespace ssp;
  if (!ctx) return SSP_ERR_INVALID;
  if (int rc = use_device(ctx)) return rc;
  if (int rc = finish_pending_measures(ctx, "ssp_polygon_upload_features")) return rc;
  if (int rc = ensure_poly_image(ctx, image, "ssp_polygon_upload_features", false)) return rc;
  SSP_REQUIRE(ctx, n_polygons >= 0 && n_polygons < (1LL << 31) - 1024, SSP_ERR_INVALID,
              "ssp_polygon_upload_features: bad polygon count");
  SSP_REQUIRE(ctx, n_polygons == 0 || feat != nullptr, SSP_ERR_INVALID,
              "ssp_polygon_upload_features: NULL features");
  PolyImage& pi = ctx->poly[(size_t)image];
  const cudaMemcpyKind kind = to_device_kind(mem_kind);
  SSP_CUDA_TRY(ctx, pi.feat.ensure((size_t)n_polygons * 120 + 16));
  if (n_polygons > 0)
    SSP_CUDA_TRY(ctx, cudaMemcpyAsync(pi.feat.p, feat, (size_t)n_polygons * 120, kind, ctx->stream));
  if (star_idx) {
    SSP_CUDA_TRY(ctx, pi.star_idx.ensure((size_t)n_polygons * 20 + 16));
    if (n_polygons > 0)
      SSP_CUDA_TRY(ctx, cudaMemcpyAsync(pi.star_idx.p, star_idx, (size_t)n_polygons * 20, kind,
                                        ctx->stream));
  } else {
    pi.star_idx.release();
  }
  if (mem_kind != SSP_MEM_DEVICE) SSP_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));  // host source may be a temporary
  pi.count = n_polygons;
  pi.has_raw = false;
  ctx->p_n_matches = -1;
  return SSP_OK;
}


int ssp_polygon_match_pairs(ssp_ctx* ctx, const int32_t* pairs, int n_pairs, double threshold,
                            int shift_mode, int precision, int64_t* match_offsets) {
  using namespace ssp;
  if (!ctx) return SSP_ERR_INVALID;
  if (int rc = use_device(ctx)) return rc;
  if (int rc = finish_pending_measures(ctx, "ssp_polygon_match_pairs")) return rc;
  SSP_REQUIRE(ctx, n_pairs >= 0 && (n_pairs == 0 || pairs != nullptr), SSP_ERR_INVALID,
              "ssp_polygon_match_pairs: bad pair list");
  SSP_REQUIRE(ctx, match_offsets != nullptr, SSP_ERR_INVALID,
              "ssp_polygon_match_pairs: match_offsets is NULL");
  SSP_REQUIRE(ctx, shift_mode == SSP_SHIFT_CANONICAL || shift_mode == SSP_SHIFT_MIN_OVER_SHIFTS,
              SSP_ERR_INVALID, "ssp_polygon_match_pairs: unknown shift mode %d", shift_mode);
  SSP_REQUIRE(ctx, precision == SSP_PRECISION_FP64 || precision == SSP_PRECISION_FP32,
              SSP_ERR_INVALID, "ssp_polygon_match_pairs: unknown precision %d", precision);
  ctx->p_n_matches = -1;
  ctx->p_launches = 0;
  ctx->p_main_ms = 0.f;
  ctx->p_evaluations = 0;
  ctx->p_match_offsets.assign((size_t)n_pairs + 1, 0);
  if (n_pairs == 0) {
    match_offsets[0] = 0;
    ctx->p_n_matches = 0;
    ctx->p_total_rows = 0;
    return SSP_OK;
  }
  std::vector<PolyPair> pp((size_t)n_pairs);
  long long rows = 0;
  int max_p1 = 0, max_p2 = 0;
  for (int p = 0; p < n_pairs; ++p) {
    const int i1 = pairs[2 * p], i2 = pairs[2 * p + 1];
    SSP_REQUIRE(ctx, i1 >= 0 && (size_t)i1 < ctx->poly.size() && ctx->poly[(size_t)i1].count >= 0 &&
                         i2 >= 0 && (size_t)i2 < ctx->poly.size() && ctx->poly[(size_t)i2].count >= 0,
                SSP_ERR_STATE, "ssp_polygon_match_pairs: pair %d = (%d, %d) has no polygon bank", p,
                i1, i2);
    pp[p].f1 = ctx->poly[(size_t)i1].feat.as<double>();
    pp[p].f2 = ctx->poly[(size_t)i2].feat.as<double>();
    pp[p].p1 = (int)ctx->poly[(size_t)i1].count;
    pp[p].p2 = (int)ctx->poly[(size_t)i2].count;
    pp[p].out_row = rows;
    rows += pp[p].p1;
    max_p1 = pp[p].p1 > max_p1 ? pp[p].p1 : max_p1;
    max_p2 = pp[p].p2 > max_p2 ? pp[p].p2 : max_p2;
    ctx->p_evaluations += (long long)pp[p].p1 * pp[p].p2;
  }
  ctx->p_total_rows = rows;
  SSP_CUDA_TRY(ctx, ctx->p_pairs.ensure(pp.size() * sizeof(PolyPair)));
  SSP_CUDA_TRY(ctx, ctx->p_argmin.ensure((size_t)(rows + 1) * 4));
  SSP_CUDA_TRY(ctx, ctx->p_mindist.ensure((size_t)(rows + 1) * 8));
  SSP_CUDA_TRY(ctx, ctx->p_counts.ensure((size_t)n_pairs * 4));
  SSP_CUDA_TRY(ctx, ctx->p_offsets.ensure(((size_t)n_pairs + 1) * 8));
  SSP_CUDA_TRY(ctx, cudaMemcpyAsync(ctx->p_pairs.p, pp.data(), pp.size() * sizeof(PolyPair),
                                    cudaMemcpyHostToDevice, ctx->stream));
  // matches: at most one per row of the first image, so the output arrays are sized up front and the
  // call synchronises with the host once, at its end
  SSP_CUDA_TRY(ctx, ctx->p_m1.ensure((size_t)(rows + 1) * 4));
  SSP_CUDA_TRY(ctx, ctx->p_m2.ensure((size_t)(rows + 1) * 4));
  SSP_CUDA_TRY(ctx, ctx->p_md.ensure((size_t)(rows + 1) * 8));
  SSP_CUDA_TRY(ctx, ctx->p_flops.ensure(16));
  ctx->p_executed_flops = -1;
  ctx->p_executed_flops_f32 = -1;
  // Which exact form.  SSP_B200_POLY_COMPARE = tensor | screen | queue | legacy picks one (A/B runs and tests);
  // by default the TF32 tensor-core screen with FP64 evaluation of its candidates (polygon_tensor.cuh), which
  // hands the call over to the CUDA-core kernels when it must (features beyond the float range, candidate
  // pool overflow, a screen value further from FP64 than its bound).  Of the CUDA-core kernels the
  // single-precision screen costs the same ~1.1 ps per cell whatever the data (52 ms at the 15 000 cap
  // against 69 ms), the first-group queue less where real matches tighten the rows' thresholds early (2.6
  // against 3.5 ms at C4, 3.4 k polygons per image).
  const char* env = getenv("SSP_B200_POLY_COMPARE");
  const char kind = env ? env[0] : 0;
  const bool ms = (shift_mode == SSP_SHIFT_MIN_OVER_SHIFTS);
  bool tensor = precision == SSP_PRECISION_FP64 && !ms && max_p1 > 0 && max_p2 > 0 && (kind == 0 || kind == 't');
  ctx->t_used = 0;
  ctx->t_fallback_flags = 0;
  ctx->t_candidates = -1;
  ctx->t_worst_coef = -1.0;
  long long total = 0;
  for (int attempt = 0; attempt < 2; ++attempt) {
    bool counted = false;
    ctx->p_launches = 0;
    SSP_CUDA_TRY(ctx, cudaMemsetAsync(ctx->p_flops.p, 0, 16, ctx->stream));
    if (max_p1 > 0 && tensor) {
      NvtxRange range("kernel 3 (polygon compare, tensor-core screen)");
      SSP_CUDA_TRY(ctx, launch_poly_compare_tensor(ctx, pp, pairs, n_pairs, rows));
      counted = true;
    } else if (max_p1 > 0) {
      NvtxRange range("kernel 3 (polygon compare)");
      dim3 grid((unsigned)n_pairs, (unsigned)((max_p1 + PC_THREADS - 1) / PC_THREADS));
      SSP_CUDA_TRY(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
      const PolyPair* dp = ctx->p_pairs.as<PolyPair>();
      int* am = ctx->p_argmin.as<int>();
      double* md = ctx->p_mindist.as<double>();
      const bool legacy = kind == 'l';
      const bool screen_fits = max_p2 < PS_MAX_COLS;
      const bool queue = kind == 'q' || !screen_fits || (kind != 's' && max_p2 < PS_AUTO_MIN_COLS);
      if (precision == SSP_PRECISION_FP64) {
        if (ms) poly_compare_f64_kernel<true><<<grid, PC_THREADS, 0, ctx->stream>>>(dp, am, md);
        else if (legacy) poly_compare_f64_kernel<false><<<grid, PC_THREADS, 0, ctx->stream>>>(dp, am, md);
        else if (queue) {
          poly_compare_f64_queue_kernel<<<grid, PC_THREADS, 0, ctx->stream>>>(
              dp, am, md, ctx->p_flops.as<unsigned long long>());
          counted = true;
        } else {
          dim3 grid_s((unsigned)n_pairs, (unsigned)((max_p1 + PS_CTA_ROWS - 1) / PS_CTA_ROWS));
          poly_compare_f64_screen_kernel<<<grid_s, PC_THREADS, 0, ctx->stream>>>(
              dp, am, md, ctx->p_flops.as<unsigned long long>());
          counted = true;
        }
      } else {
        if (ms) poly_compare_f32_kernel<true><<<grid, PC_THREADS, 0, ctx->stream>>>(dp, am, md);
        else poly_compare_f32_kernel<false><<<grid, PC_THREADS, 0, ctx->stream>>>(dp, am, md);
      }
      SSP_CUDA_TRY(ctx, cudaGetLastError());
      SSP_CUDA_TRY(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
      ctx->p_launches += 1;
    }
    {
      NvtxRange range("finalize (count, scan, compact)");
      poly_count_kernel<<<n_pairs, PF_THREADS, 0, ctx->stream>>>(
          ctx->p_pairs.as<PolyPair>(), ctx->p_mindist.as<double>(), threshold, ctx->p_counts.as<int>());
      SSP_CUDA_TRY(ctx, cudaGetLastError());
      scan_counts_kernel<<<1, 1024, 0, ctx->stream>>>(ctx->p_counts.as<int>(), n_pairs,
                                                      ctx->p_offsets.as<long long>());
      SSP_CUDA_TRY(ctx, cudaGetLastError());
      poly_compact_kernel<<<n_pairs, PF_THREADS, 0, ctx->stream>>>(
          ctx->p_pairs.as<PolyPair>(), ctx->p_argmin.as<int>(), ctx->p_mindist.as<double>(), threshold,
          ctx->p_offsets.as<long long>(), ctx->p_m1.as<int>(), ctx->p_m2.as<int>(),
          ctx->p_md.as<double>());
      SSP_CUDA_TRY(ctx, cudaGetLastError());
      ctx->p_launches += 3;
    }
    unsigned long long flops[2] = {0, 0};
    int t_flags = 0;
    unsigned t_worst = 0;
    std::vector<unsigned> t_counts;
    SSP_CUDA_TRY(ctx, cudaMemcpyAsync(ctx->p_match_offsets.data(), ctx->p_offsets.p,
                                      ((size_t)n_pairs + 1) * 8, cudaMemcpyDeviceToHost, ctx->stream));
    SSP_CUDA_TRY(ctx, cudaMemcpyAsync(flops, ctx->p_flops.p, 16, cudaMemcpyDeviceToHost, ctx->stream));
    if (tensor && max_p1 > 0) {
      t_counts.assign((size_t)ctx->t_grid, 0u);
      SSP_CUDA_TRY(ctx, cudaMemcpyAsync(&t_flags, ctx->t_flags.p, 4, cudaMemcpyDeviceToHost, ctx->stream));
      SSP_CUDA_TRY(ctx, cudaMemcpyAsync(&t_worst, ctx->t_f.as<unsigned>() + ctx->t_stats_off, 4,
                                        cudaMemcpyDeviceToHost, ctx->stream));
      SSP_CUDA_TRY(ctx, cudaMemcpyAsync(t_counts.data(), ctx->t_counts.p, t_counts.size() * 4,
                                        cudaMemcpyDeviceToHost, ctx->stream));
    }
    SSP_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    if (tensor && max_p1 > 0) {
      float worst;
      memcpy(&worst, &t_worst, 4);
      ctx->t_worst_coef = (double)worst;
      ctx->t_candidates = 0;
      for (unsigned c : t_counts) ctx->t_candidates += c;
      if (t_flags != 0) {          // hand the call over to the CUDA-core kernels
        ctx->t_fallback_flags = t_flags;
        tensor = false;
        continue;
      }
      ctx->t_used = 1;
    }
    total = ctx->p_match_offsets[(size_t)n_pairs];
    if (counted) {
      ctx->p_executed_flops = (long long)(flops[0] + flops[1]);
      ctx->p_executed_flops_f32 = (long long)flops[1];
    }
    if (max_p1 > 0) SSP_CUDA_TRY(ctx, cudaEventElapsedTime(&ctx->p_main_ms, ctx->ev0, ctx->ev1));
    break;
  }
  for (int p = 0; p <= n_pairs; ++p) match_offsets[p] = ctx->p_match_offsets[(size_t)p];
  ctx->p_n_matches = total;
  return SSP_OK;
}

int ssp_polygon_fetch_matches(ssp_ctx* ctx, int32_t* poly1, int32_t* poly2, double* dist,
                              int64_t capacity, int mem_kind) {
  using namespace ssp;
  if (!ctx) return SSP_ERR_INVALID;
  if (int rc = use_device(ctx)) return rc;
  SSP_REQUIRE(ctx, ctx->p_n_matches >= 0, SSP_ERR_STATE,
              "ssp_polygon_fetch_matches: no completed ssp_polygon_match_pairs call");
  const long long n = ctx->p_n_matches;
  SSP_REQUIRE(ctx, capacity >= n, SSP_ERR_CAPACITY, "ssp_polygon_fetch_matches: capacity too small");
  if (n == 0) return SSP_OK;
  const cudaMemcpyKind kind = from_device_kind(mem_kind);
  if (poly1) SSP_CUDA_TRY(ctx, cudaMemcpyAsync(poly1, ctx->p_m1.p, (size_t)n * 4, kind, ctx->stream));
  if (poly2) SSP_CUDA_TRY(ctx, cudaMemcpyAsync(poly2, ctx->p_m2.p, (size_t)n * 4, kind, ctx->stream));
  if (dist) SSP_CUDA_TRY(ctx, cudaMemcpyAsync(dist, ctx->p_md.p, (size_t)n * 8, kind, ctx->stream));
  SSP_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  return SSP_OK;
}

int ssp_polygon_fetch_rowmin(ssp_ctx* ctx, int32_t* argmin, double* mindist, int64_t capacity_rows,
                             int mem_kind) {
  using namespace ssp;
  if (!ctx) return SSP_ERR_INVALID;
  if (int rc = use_device(ctx)) return rc;
  SSP_REQUIRE(ctx, ctx->p_n_matches >= 0, SSP_ERR_STATE,
              "ssp_polygon_fetch_rowmin: no completed ssp_polygon_match_pairs call");
  const long long n = ctx->p_total_rows;
  SSP_REQUIRE(ctx, capacity_rows >= n, SSP_ERR_CAPACITY, "ssp_polygon_fetch_rowmin: capacity too small");
  if (n == 0) return SSP_OK;
  const cudaMemcpyKind kind = from_device_kind(mem_kind);
  if (argmin) SSP_CUDA_TRY(ctx, cudaMemcpyAsync(argmin, ctx->p_argmin.p, (size_t)n * 4, kind, ctx->stream));
  if (mindist) SSP_CUDA_TRY(ctx, cudaMemcpyAsync(mindist, ctx->p_mindist.p, (size_t)n * 8, kind, ctx->stream));
  SSP_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  return SSP_OK;
}

int ssp_polygon_last_screen_stats(ssp_ctx* ctx, int* used_tensor, int* fallback_flags, int64_t* candidates,
                                  double* worst_error_coef, double* bound_coef) {
  if (!ctx || !used_tensor || !fallback_flags || !candidates || !worst_error_coef || !bound_coef) return SSP_ERR_INVALID;
  *used_tensor = ctx->t_used;
  *fallback_flags = ctx->t_fallback_flags;
  *candidates = ctx->t_candidates;
  *worst_error_coef = ctx->t_worst_coef;
  *bound_coef = (double)ssp::PT_ERR_COEF;
  return SSP_OK;
}

int ssp_polygon_last_executed_flops_by_precision(ssp_ctx* ctx, int64_t* fp64_flops, int64_t* fp32_flops) {
  if (!ctx || !fp64_flops || !fp32_flops) return SSP_ERR_INVALID;
  *fp64_flops = ctx->p_executed_flops < 0 ? -1 : ctx->p_executed_flops - ctx->p_executed_flops_f32;
  *fp32_flops = ctx->p_executed_flops < 0 ? -1 : ctx->p_executed_flops_f32;
  return SSP_OK;
}

int ssp_polygon_last_executed_flops(ssp_ctx* ctx, int64_t* executed_flops) {
  if (!ctx || !executed_flops) return SSP_ERR_INVALID;
  *executed_flops = ctx->p_executed_flops;
  return SSP_OK;
}

int ssp_polygon_last_stats(ssp_ctx* ctx, int* n_launches, float* main_kernel_ms,
                           int64_t* evaluations) {
  if (!ctx) return SSP_ERR_INVALID;
  if (n_launches) *n_launches = ctx->p_launches;
  if (main_kernel_ms) *main_kernel_ms = ctx->p_main_ms;
  if (evaluations) *evaluations = ctx->p_evaluations;
  return SSP_OK;
}

}  // extern "C"
