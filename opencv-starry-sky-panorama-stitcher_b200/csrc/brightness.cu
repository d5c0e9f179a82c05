// Row N1 — get_star_brightness (image_processors.py:222-254) for all stars of one image.
//
// The reference masks the full grey frame with a filled radius-10 circle and sorts all 1.2 M
// pixels for every star (90 % of its polygon-path wall time).  Everything outside the disc is
// zero after the mask, so the result is the sum of the `top_k` largest grey values among the
// disc pixels that fall inside the image.  One warp per star: each lane keeps its share of the
// disc in registers, then top_k rounds of (lane max -> warp max -> retire one instance).
#include "common.cuh"

namespace ssp {

constexpr int SB_MAX_PER_LANE = 16;  // supports discs up to 512 pixels (radius 10 -> 317)

__global__ void __launch_bounds__(128)
star_brightness_kernel(const unsigned char* __restrict__ gray, int h, int w,
                       const int* __restrict__ center_xy, long long n,
                       const int* __restrict__ disc, int n_disc, int top_k,
                       int* __restrict__ out) {
  const long long star = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (star >= n) return;
  const int cx = center_xy[2 * star], cy = center_xy[2 * star + 1];
  int v[SB_MAX_PER_LANE];
#pragma unroll
  for (int k = 0; k < SB_MAX_PER_LANE; ++k) {
    const int i = k * 32 + lane;
    int val = -1;  // -1: no pixel (outside the disc list or outside the image)
    if (i < n_disc) {
      const int x = cx + disc[2 * i], y = cy + disc[2 * i + 1];
      if (x >= 0 && x < w && y >= 0 && y < h) val = gray[(long long)y * w + x];
    }
    v[k] = val;
  }
  int sum = 0;
  for (int round = 0; round < top_k; ++round) {
    int m = -1, mk = 0;
#pragma unroll
    for (int k = 0; k < SB_MAX_PER_LANE; ++k)
      if (v[k] > m) {
        m = v[k];
        mk = k;
      }
    int wm = m;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) wm = max(wm, __shfl_xor_sync(0xffffffffu, wm, o));
    if (wm < 0) break;  // fewer than top_k pixels: the remaining sorted entries are zeros
    sum += wm;
    const unsigned who = __ballot_sync(0xffffffffu, m == wm);
    if (lane == __ffs(who) - 1) {
#pragma unroll
      for (int k = 0; k < SB_MAX_PER_LANE; ++k)
        if (k == mk) v[k] = -1;
    }
  }
  if (lane == 0) out[star] = sum;
}

}  // namespace ssp

extern "C" int ssp_star_brightness(ssp_ctx* ctx, const uint8_t* gray, int h, int w,
                                   const int32_t* center_xy, int64_t n, const int32_t* disc_dxdy,
                                   int n_disc, int top_k, int32_t* brightness, int mem_kind) {
  using namespace ssp;
  if (!ctx) return SSP_ERR_INVALID;
  if (int rc = use_device(ctx)) return rc;
  SSP_REQUIRE(ctx, h > 0 && w > 0 && gray != nullptr, SSP_ERR_INVALID,
              "ssp_star_brightness: bad image");
  SSP_REQUIRE(ctx, n >= 0 && (n == 0 || (center_xy != nullptr && brightness != nullptr)),
              SSP_ERR_INVALID, "ssp_star_brightness: bad star arrays");
  SSP_REQUIRE(ctx, n_disc > 0 && n_disc <= 32 * SB_MAX_PER_LANE && disc_dxdy != nullptr,
              SSP_ERR_INVALID, "ssp_star_brightness: disc must have 1..%d pixels",
              32 * SB_MAX_PER_LANE);
  SSP_REQUIRE(ctx, top_k > 0 && top_k <= n_disc, SSP_ERR_INVALID, "ssp_star_brightness: bad top_k");
  if (n == 0) return SSP_OK;
  const cudaMemcpyKind in_kind = to_device_kind(mem_kind);
  const size_t img_bytes = (size_t)h * w;
  SSP_CUDA_TRY(ctx, ctx->b_img.ensure(img_bytes));
  SSP_CUDA_TRY(ctx, ctx->b_xy.ensure((size_t)n * 8));
  SSP_CUDA_TRY(ctx, ctx->b_disc.ensure((size_t)n_disc * 8));
  SSP_CUDA_TRY(ctx, ctx->b_out.ensure((size_t)n * 4));
  SSP_CUDA_TRY(ctx, cudaMemcpyAsync(ctx->b_img.p, gray, img_bytes, in_kind, ctx->stream));
  SSP_CUDA_TRY(ctx, cudaMemcpyAsync(ctx->b_xy.p, center_xy, (size_t)n * 8, in_kind, ctx->stream));
  // the disc offsets are always a small host array
  SSP_CUDA_TRY(ctx, cudaMemcpyAsync(ctx->b_disc.p, disc_dxdy, (size_t)n_disc * 8,
                                    cudaMemcpyHostToDevice, ctx->stream));
  const long long threads = n * 32;
  star_brightness_kernel<<<(unsigned)((threads + 127) / 128), 128, 0, ctx->stream>>>(
      ctx->b_img.as<unsigned char>(), h, w, ctx->b_xy.as<int>(), n, ctx->b_disc.as<int>(), n_disc,
      top_k, ctx->b_out.as<int>());
  SSP_CUDA_TRY(ctx, cudaGetLastError());
  SSP_CUDA_TRY(ctx, cudaMemcpyAsync(brightness, ctx->b_out.p, (size_t)n * 4,
                                    from_device_kind(mem_kind), ctx->stream));
  SSP_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  return SSP_OK;
}
