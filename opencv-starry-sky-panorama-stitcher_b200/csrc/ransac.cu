// Row N3 — batched RANSAC homographies for all image pairs of the last Hamming call (sm_100a).
//
// Stands in, opt-in, for the per-pair cv.findHomography(src_pts, dst_pts, cv.RANSAC, 5) of
// custom_bruteforce_matcher.py:215-246 (and :444-475 on the star path): the step right behind
// kernel 1, which at the reference's real sizes (<= 1000 descriptors per image) is where the
// whole matcher stage spends its time (59 ms per pair on the host against microseconds of
// matching).  The survivors of ssp_hamming_match_pairs are still on the device, so the kernel
// gathers its point pairs straight from them: no host round trip between matching and RANSAC.
//
// NOT bit-compatible with OpenCV by nature (OpenCV's RNG, adaptive iteration count and final
// Levenberg-Marquardt polish are its own): same model (4-point DLT hypotheses, reprojection
// error <= threshold counts as inlier, best hypothesis wins, then least-squares refits on the inliers
// that are accepted while they gain inliers), fixed seed, fixed hypothesis count.  tests/ and bench.py report inlier-count agreement with cv.findHomography.
//
// One CTA per pair.  Points are Hartley-normalised per pair; every thread owns hypotheses
// tid, tid + 256, ...: counter-based sampling of 4 distinct matches, degeneracy test (collinear
// triples / orientation flips, as cv's checkSubset), 8 x 8 Gaussian elimination with partial
// pivoting in float64, inlier count over the point pairs staged in shared memory.
#include "common.cuh"

namespace ssp {

constexpr int RS_THREADS = 256;
constexpr int RS_TILE = 1024;   // point pairs per shared-memory tile

struct RansacPair {
  long long match_off;   // first survivor of the pair in the engine's match arrays
  int n;                 // number of matches
  long long kp_q, kp_t;  // first keypoint of the query / train image in the keypoint table
};

__device__ __forceinline__ uint32_t rs_hash(uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  // PCG-style mixing of four counters: the k-th index of hypothesis h of pair p, attempt a
  uint32_t x = a * 0x9E3779B1u ^ (b + 0x7F4A7C15u) * 0x85EBCA77u ^ (c + 0x165667B1u) * 0xC2B2AE3Du ^ d * 0x27D4EB2Fu;
  x ^= x >> 15; x *= 0x2C1B3C6Du; x ^= x >> 12; x *= 0x297A2D39u; x ^= x >> 15;
  return x;
}

__device__ __forceinline__ double rs_cross(double ax, double ay, double bx, double by, double cx, double cy) {
  return (bx - ax) * (cy - ay) - (by - ay) * (cx - ax);
}

// h (8 unknowns, h33 = 1) of the homography through 4 correspondences (x, y) -> (u, v); false if singular
__device__ bool rs_solve4(const double (&x)[4], const double (&y)[4], const double (&u)[4], const double (&v)[4],
                          double (&h)[9]) {
  double A[8][9];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    double* r0 = A[2 * i];
    double* r1 = A[2 * i + 1];
    r0[0] = x[i]; r0[1] = y[i]; r0[2] = 1.0; r0[3] = 0.0; r0[4] = 0.0; r0[5] = 0.0; r0[6] = -x[i] * u[i]; r0[7] = -y[i] * u[i]; r0[8] = u[i];
    r1[0] = 0.0; r1[1] = 0.0; r1[2] = 0.0; r1[3] = x[i]; r1[4] = y[i]; r1[5] = 1.0; r1[6] = -x[i] * v[i]; r1[7] = -y[i] * v[i]; r1[8] = v[i];
  }
  for (int c = 0; c < 8; ++c) {
    int piv = c;
    double best = fabs(A[c][c]);
    for (int r = c + 1; r < 8; ++r)
      if (fabs(A[r][c]) > best) { best = fabs(A[r][c]); piv = r; }
    if (best < 1e-12) return false;
    if (piv != c)
      for (int k = c; k < 9; ++k) { const double t = A[c][k]; A[c][k] = A[piv][k]; A[piv][k] = t; }
    const double inv = 1.0 / A[c][c];
    for (int r = c + 1; r < 8; ++r) {
      const double f = A[r][c] * inv;
      if (f != 0.0)
        for (int k = c; k < 9; ++k) A[r][k] -= f * A[c][k];
    }
  }
  for (int c = 7; c >= 0; --c) {
    double s = A[c][8];
    for (int k = c + 1; k < 8; ++k) s -= A[c][k] * h[k];
    h[c] = s / A[c][c];
  }
  h[8] = 1.0;
  return true;
}

// H_pix = Td^-1 * Hn * Ts for similarity normalisations T(p) = s * (p - c)
__device__ __forceinline__ void rs_denormalise(const double (&hn)[9], double cs_x, double cs_y, double ss,
                                               double cd_x, double cd_y, double sd, double (&hp)[9]) {
  // Hn * Ts
  double m[9];
#pragma unroll
  for (int r = 0; r < 3; ++r) {
    m[3 * r] = hn[3 * r] * ss;
    m[3 * r + 1] = hn[3 * r + 1] * ss;
    m[3 * r + 2] = hn[3 * r + 2] - ss * (hn[3 * r] * cs_x + hn[3 * r + 1] * cs_y);
  }
  // Td^-1 = [[1/sd, 0, cd_x], [0, 1/sd, cd_y], [0, 0, 1]]
  const double isd = 1.0 / sd;
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    hp[k] = m[k] * isd + cd_x * m[6 + k];
    hp[3 + k] = m[3 + k] * isd + cd_y * m[6 + k];
    hp[6 + k] = m[6 + k];
  }
}

__global__ void __launch_bounds__(RS_THREADS)
ransac_homography_kernel(const RansacPair* __restrict__ pairs, const int* __restrict__ mq,
                         const int* __restrict__ mt, const float* __restrict__ kp_xy, int n_hyp,
                         float thresh_sq, uint32_t seed, int min_matches, double* __restrict__ H_out,
                         unsigned char* __restrict__ mask_out, int* __restrict__ n_inliers_out) {
  __shared__ float4 pts[RS_TILE];                 // (x, y, u, v) of one tile of point pairs
  __shared__ double red[RS_THREADS];
  __shared__ unsigned long long best_key_s;
  __shared__ double norm_s[6];                    // cs_x, cs_y, ss, cd_x, cd_y, sd
  __shared__ float best_h[9];
  __shared__ double ata[RS_THREADS / 32][44];
  const RansacPair pr = pairs[blockIdx.x];
  const int n = pr.n;
  const int tid = threadIdx.x;
  double* Hp = H_out + 9 * (long long)blockIdx.x;
  if (n < min_matches) {   // custom_bruteforce_matcher.py:215: fewer than 6 matches -> no homography
    if (tid < 9) Hp[tid] = nan("");
    if (tid == 0) n_inliers_out[blockIdx.x] = -1;
    return;
  }
  auto load_pt = [&](int i) {
    const int q = mq[pr.match_off + i], t = mt[pr.match_off + i];
    return make_float4(kp_xy[2 * (pr.kp_q + q)], kp_xy[2 * (pr.kp_q + q) + 1], kp_xy[2 * (pr.kp_t + t)],
                       kp_xy[2 * (pr.kp_t + t) + 1]);
  };
  auto block_sum = [&](double v) {
    red[tid] = v;
    __syncthreads();
    for (int o = RS_THREADS / 2; o > 0; o >>= 1) {
      if (tid < o) red[tid] += red[tid + o];
      __syncthreads();
    }
    const double r = red[0];
    __syncthreads();
    return r;
  };
  // ---- Hartley normalisation of both point sets --------------------------------------------------
  double sx = 0, sy = 0, su = 0, sv = 0;
  for (int i = tid; i < n; i += RS_THREADS) {
    const float4 p = load_pt(i);
    sx += p.x; sy += p.y; su += p.z; sv += p.w;
  }
  const double cs_x = block_sum(sx) / n, cs_y = block_sum(sy) / n, cd_x = block_sum(su) / n, cd_y = block_sum(sv) / n;
  double ds = 0, dd = 0;
  for (int i = tid; i < n; i += RS_THREADS) {
    const float4 p = load_pt(i);
    ds += sqrt((p.x - cs_x) * (p.x - cs_x) + (p.y - cs_y) * (p.y - cs_y));
    dd += sqrt((p.z - cd_x) * (p.z - cd_x) + (p.w - cd_y) * (p.w - cd_y));
  }
  const double ms = block_sum(ds) / n, md = block_sum(dd) / n;
  const double ss = ms > 1e-9 ? 1.4142135623730951 / ms : 1.0, sd = md > 1e-9 ? 1.4142135623730951 / md : 1.0;
  if (tid == 0) {
    norm_s[0] = cs_x; norm_s[1] = cs_y; norm_s[2] = ss; norm_s[3] = cd_x; norm_s[4] = cd_y; norm_s[5] = sd;
    best_key_s = 0ull;
  }
  __syncthreads();

  // ---- hypotheses: tid, tid + 256, ... --------------------------------------------------------------
  for (int h0 = 0; h0 < n_hyp; h0 += RS_THREADS) {
    const int hyp = h0 + tid;
    float hf[9];
    bool valid = hyp < n_hyp;
    if (valid) {
      int idx[4];
      uint32_t attempt = 0;
      for (int k = 0; k < 4 && valid; ++k) {
        bool fresh = false;
        for (int tries = 0; tries < 16 && !fresh; ++tries) {
          idx[k] = (int)(rs_hash(seed, blockIdx.x, (uint32_t)hyp, attempt++) % (uint32_t)n);
          fresh = true;
          for (int j = 0; j < k; ++j) fresh = fresh && idx[j] != idx[k];
        }
        valid = fresh;
      }
      double x[4], y[4], u[4], v[4];
      if (valid) {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const float4 p = load_pt(idx[k]);
          x[k] = (p.x - cs_x) * ss; y[k] = (p.y - cs_y) * ss; u[k] = (p.z - cd_x) * sd; v[k] = (p.w - cd_y) * sd;
        }
        // no three points collinear, and every triple keeps its orientation (cv::checkSubset)
#pragma unroll
        for (int a = 0; a < 2; ++a)
#pragma unroll
          for (int b = a + 1; b < 3; ++b)
#pragma unroll
            for (int c = b + 1; c < 4; ++c) {
              const double o1 = rs_cross(x[a], y[a], x[b], y[b], x[c], y[c]);
              const double o2 = rs_cross(u[a], v[a], u[b], v[b], u[c], v[c]);
              valid = valid && fabs(o1) > 1e-6 && fabs(o2) > 1e-6 && (o1 > 0) == (o2 > 0);
            }
      }
      double hn[9], hp[9];
      if (valid) valid = rs_solve4(x, y, u, v, hn);
      if (valid) {
        rs_denormalise(hn, cs_x, cs_y, ss, cd_x, cd_y, sd, hp);
        valid = fabs(hp[8]) > 1e-12;
        if (valid) {
          const double inv = 1.0 / hp[8];
#pragma unroll
          for (int k = 0; k < 9; ++k) hf[k] = (float)(hp[k] * inv);
        }
      }
    }
    if (!valid) {
#pragma unroll
      for (int k = 0; k < 9; ++k) hf[k] = 0.f;
    }
    // inliers of my hypothesis over all point pairs, staged tile by tile
    int count = 0;
    for (int base = 0; base < n; base += RS_TILE) {
      const int m = min(RS_TILE, n - base);
      __syncthreads();
      for (int i = tid; i < m; i += RS_THREADS) pts[i] = load_pt(base + i);
      __syncthreads();
      if (valid) {
        for (int i = 0; i < m; ++i) {
          const float4 p = pts[i];
          const float w = hf[6] * p.x + hf[7] * p.y + hf[8];
          const float iw = 1.f / w;
          const float dx = (hf[0] * p.x + hf[1] * p.y + hf[2]) * iw - p.z;
          const float dy = (hf[3] * p.x + hf[4] * p.y + hf[5]) * iw - p.w;
          count += (dx * dx + dy * dy <= thresh_sq) ? 1 : 0;
        }
      }
    }
    // best = most inliers, then the lowest hypothesis number (deterministic)
    const unsigned long long key = valid ? (((unsigned long long)count << 32) | (unsigned)(0x7FFFFFFF - hyp)) : 0ull;
    const unsigned long long old = atomicMax(&best_key_s, key);
    __syncthreads();
    if (valid && key > old && best_key_s == key) {
#pragma unroll
      for (int k = 0; k < 9; ++k) best_h[k] = hf[k];
    }
    __syncthreads();
  }
  const int best_count = (int)(best_key_s >> 32);
  if (best_count < 4) {   // no usable model
    if (tid < 9) Hp[tid] = nan("");
    for (int i = tid; i < n; i += RS_THREADS) mask_out[pr.match_off + i] = 0;
    if (tid == 0) n_inliers_out[blockIdx.x] = 0;
    return;
  }
  // ---- local optimisation: least-squares refit (h33 = 1, normalised) on the inliers of the current
  // model, accepted while it gains inliers (a 4-point model through noisy points rarely holds all
  // of them; cv::findHomography polishes with Levenberg-Marquardt instead) ----------------------------
  __shared__ float cur_h[9];
  __shared__ double refit_h[9];
  __shared__ int refit_ok;
  if (tid < 9) cur_h[tid] = best_h[tid];
  __syncthreads();
  auto is_inlier = [&](const float (&h)[9], const float4 p) {
    const float w = h[6] * p.x + h[7] * p.y + h[8];
    const float iw = 1.f / w;
    const float dx = (h[0] * p.x + h[1] * p.y + h[2]) * iw - p.z;
    const float dy = (h[3] * p.x + h[4] * p.y + h[5]) * iw - p.w;
    return dx * dx + dy * dy <= thresh_sq;
  };
  // least squares over the inliers of cur_h -> refit_h (pixel space, h33 = 1), refit_ok
  auto refit = [&]() {
    float hb[9];
#pragma unroll
    for (int k = 0; k < 9; ++k) hb[k] = cur_h[k];
    double acc[44];
#pragma unroll
    for (int k = 0; k < 44; ++k) acc[k] = 0.0;
    for (int i = tid; i < n; i += RS_THREADS) {
      const float4 p = load_pt(i);
      if (!is_inlier(hb, p)) continue;
      const double x = (p.x - cs_x) * ss, y = (p.y - cs_y) * ss, u = (p.z - cd_x) * sd, v = (p.w - cd_y) * sd;
      const double r0[9] = {x, y, 1.0, 0.0, 0.0, 0.0, -x * u, -y * u, u};
      const double r1[9] = {0.0, 0.0, 0.0, x, y, 1.0, -x * v, -y * v, v};
      // upper triangle of A^T A (8 x 8: 36 entries) and A^T b (8 entries)
      int e = 0;
#pragma unroll
      for (int a = 0; a < 8; ++a)
#pragma unroll
        for (int b = a; b < 8; ++b) acc[e++] += r0[a] * r0[b] + r1[a] * r1[b];
#pragma unroll
      for (int a = 0; a < 8; ++a) acc[36 + a] += r0[a] * r0[8] + r1[a] * r1[8];
    }
#pragma unroll
    for (int k = 0; k < 44; ++k) {
      double v = acc[k];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      if ((tid & 31) == 0) ata[tid >> 5][k] = v;
    }
    __syncthreads();
    if (tid == 0) {
      double M[8][9];
      int e = 0;
      for (int a = 0; a < 8; ++a)
        for (int b = a; b < 8; ++b) {
          double s = 0.0;
          for (int w = 0; w < RS_THREADS / 32; ++w) s += ata[w][e];
          M[a][b] = s;
          M[b][a] = s;
          ++e;
        }
      for (int a = 0; a < 8; ++a) {
        double s = 0.0;
        for (int w = 0; w < RS_THREADS / 32; ++w) s += ata[w][36 + a];
        M[a][8] = s;
      }
      double hn[9];
      bool ok = true;
      for (int c = 0; c < 8 && ok; ++c) {
        int piv = c;
        double best = fabs(M[c][c]);
        for (int r = c + 1; r < 8; ++r)
          if (fabs(M[r][c]) > best) { best = fabs(M[r][c]); piv = r; }
        ok = best > 1e-14;
        if (!ok) break;
        if (piv != c)
          for (int k = c; k < 9; ++k) { const double t = M[c][k]; M[c][k] = M[piv][k]; M[piv][k] = t; }
        for (int r = c + 1; r < 8; ++r) {
          const double f = M[r][c] / M[c][c];
          for (int k = c; k < 9; ++k) M[r][k] -= f * M[c][k];
        }
      }
      if (ok) {
        for (int c = 7; c >= 0; --c) {
          double s = M[c][8];
          for (int k = c + 1; k < 8; ++k) s -= M[c][k] * hn[k];
          hn[c] = s / M[c][c];
        }
        hn[8] = 1.0;
        double hp[9];
        rs_denormalise(hn, norm_s[0], norm_s[1], norm_s[2], norm_s[3], norm_s[4], norm_s[5], hp);
        ok = fabs(hp[8]) > 1e-12;
        if (ok)
          for (int k = 0; k < 9; ++k) refit_h[k] = hp[k] / hp[8];
      }
      refit_ok = ok ? 1 : 0;
    }
    __syncthreads();
  };
  auto count_inliers = [&](const float (&h)[9]) {
    int c = 0;
    for (int i = tid; i < n; i += RS_THREADS) c += is_inlier(h, load_pt(i)) ? 1 : 0;
    return (int)(block_sum((double)c) + 0.5);
  };
  int cur_count = best_count;
  for (int round = 0; round < 4; ++round) {
    refit();
    if (!refit_ok) break;
    float hr[9];
#pragma unroll
    for (int k = 0; k < 9; ++k) hr[k] = (float)refit_h[k];
    const int c = count_inliers(hr);          // uniform over the block
    if (c <= cur_count) break;
    __syncthreads();
    if (tid < 9) cur_h[tid] = hr[tid];
    cur_count = c;
    __syncthreads();
  }
  // the answer: mask = inliers of the accepted model, H = least squares over exactly those inliers
  refit();
  float hc[9];
#pragma unroll
  for (int k = 0; k < 9; ++k) hc[k] = cur_h[k];
  for (int i = tid; i < n; i += RS_THREADS) mask_out[pr.match_off + i] = is_inlier(hc, load_pt(i)) ? 1 : 0;
  if (tid < 9) Hp[tid] = refit_ok ? refit_h[tid] : (double)hc[tid];
  if (tid == 0) n_inliers_out[blockIdx.x] = cur_count;
}

}  // namespace ssp

extern "C" {

int ssp_ransac_homography_pairs(ssp_ctx* ctx, const float* keypoint_xy, const int64_t* keypoint_offsets,
                                int n_images, const int32_t* pairs, int n_pairs, double reproj_threshold,
                                int n_hypotheses, uint32_t seed, int min_matches, double* H,
                                uint8_t* inlier_mask, int64_t mask_capacity, int32_t* n_inliers) {
  using namespace ssp;
  if (!ctx) return SSP_ERR_INVALID;
  if (int rc = use_device(ctx)) return rc;
  SSP_REQUIRE(ctx, ctx->h_n_matches >= 0, SSP_ERR_STATE,
              "ssp_ransac_homography_pairs: no completed ssp_hamming_match_pairs call");
  SSP_REQUIRE(ctx, n_pairs == (int)ctx->h_match_offsets.size() - 1, SSP_ERR_STATE,
              "ssp_ransac_homography_pairs: %d pairs given, the last Hamming call matched %d", n_pairs,
              (int)ctx->h_match_offsets.size() - 1);
  SSP_REQUIRE(ctx, keypoint_offsets != nullptr && pairs != nullptr && H != nullptr && n_inliers != nullptr &&
                       n_images > 0,
              SSP_ERR_INVALID, "ssp_ransac_homography_pairs: NULL argument");
  SSP_REQUIRE(ctx, reproj_threshold > 0.0 && n_hypotheses > 0 && n_hypotheses <= (1 << 20) && min_matches >= 4,
              SSP_ERR_INVALID, "ssp_ransac_homography_pairs: bad threshold / hypothesis count");
  const long long total = ctx->h_n_matches;
  SSP_REQUIRE(ctx, mask_capacity >= total && (total == 0 || inlier_mask != nullptr), SSP_ERR_CAPACITY,
              "ssp_ransac_homography_pairs: mask capacity %lld < %lld matches", (long long)mask_capacity, total);
  if (n_pairs == 0) return SSP_OK;
  const long long n_kp = keypoint_offsets[n_images];
  SSP_REQUIRE(ctx, n_kp == 0 || keypoint_xy != nullptr, SSP_ERR_INVALID, "ssp_ransac_homography_pairs: NULL keypoints");
  std::vector<RansacPair> rp((size_t)n_pairs);
  for (int p = 0; p < n_pairs; ++p) {
    const int qi = pairs[2 * p], ti = pairs[2 * p + 1];
    SSP_REQUIRE(ctx, qi >= 0 && qi < n_images && ti >= 0 && ti < n_images, SSP_ERR_INVALID,
                "ssp_ransac_homography_pairs: pair %d out of range", p);
    rp[(size_t)p].match_off = ctx->h_match_offsets[(size_t)p];
    rp[(size_t)p].n = (int)(ctx->h_match_offsets[(size_t)p + 1] - ctx->h_match_offsets[(size_t)p]);
    rp[(size_t)p].kp_q = keypoint_offsets[qi];
    rp[(size_t)p].kp_t = keypoint_offsets[ti];
    // match indices are rows of the descriptor bank: they must exist in the keypoint table
    SSP_REQUIRE(ctx, keypoint_offsets[qi + 1] - keypoint_offsets[qi] >= ctx->row_off[(size_t)qi + 1] - ctx->row_off[(size_t)qi] &&
                         keypoint_offsets[ti + 1] - keypoint_offsets[ti] >= ctx->row_off[(size_t)ti + 1] - ctx->row_off[(size_t)ti],
                SSP_ERR_INVALID, "ssp_ransac_homography_pairs: pair %d: fewer keypoints than descriptor rows", p);
  }
  NvtxRange range("batched RANSAC homography");
  SSP_CUDA_TRY(ctx, ctx->r_pairs.ensure(rp.size() * sizeof(RansacPair)));
  SSP_CUDA_TRY(ctx, ctx->r_kp.ensure((size_t)n_kp * 8 + 16));
  SSP_CUDA_TRY(ctx, ctx->r_out.ensure((size_t)n_pairs * (72 + 4) + (size_t)total + 64));
  double* d_H = ctx->r_out.as<double>();
  int* d_inl = reinterpret_cast<int*>(d_H + 9 * (size_t)n_pairs);
  unsigned char* d_mask = reinterpret_cast<unsigned char*>(d_inl + n_pairs);
  SSP_CUDA_TRY(ctx, cudaMemcpyAsync(ctx->r_pairs.p, rp.data(), rp.size() * sizeof(RansacPair),
                                    cudaMemcpyHostToDevice, ctx->stream));
  if (n_kp > 0)
    SSP_CUDA_TRY(ctx, cudaMemcpyAsync(ctx->r_kp.p, keypoint_xy, (size_t)n_kp * 8, cudaMemcpyHostToDevice, ctx->stream));
  ransac_homography_kernel<<<(unsigned)n_pairs, RS_THREADS, 0, ctx->stream>>>(
      ctx->r_pairs.as<RansacPair>(), ctx->h_mq.as<int>(), ctx->h_mt.as<int>(), ctx->r_kp.as<float>(), n_hypotheses,
      (float)(reproj_threshold * reproj_threshold), seed, min_matches, d_H, d_mask, d_inl);
  SSP_CUDA_TRY(ctx, cudaGetLastError());
  SSP_CUDA_TRY(ctx, cudaMemcpyAsync(H, d_H, (size_t)n_pairs * 72, cudaMemcpyDeviceToHost, ctx->stream));
  SSP_CUDA_TRY(ctx, cudaMemcpyAsync(n_inliers, d_inl, (size_t)n_pairs * 4, cudaMemcpyDeviceToHost, ctx->stream));
  if (total > 0)
    SSP_CUDA_TRY(ctx, cudaMemcpyAsync(inlier_mask, d_mask, (size_t)total, cudaMemcpyDeviceToHost, ctx->stream));
  SSP_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  return SSP_OK;
}

}  // extern "C"
