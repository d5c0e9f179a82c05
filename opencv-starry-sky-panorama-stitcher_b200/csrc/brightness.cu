// Row N1 — get_star_brightness (image_processors.py:222-254) for all stars of one image.
//
// The reference masks the full grey frame with a filled radius-10 circle and sorts all 1.2 M
// pixels for every star (90 % of its polygon-path wall time).  Everything outside the disc is
// zero after the mask, so the result is the sum of the `top_k` largest grey values among the
// disc pixels that fall inside the image.  One warp per star: each lane keeps its share of the
// disc in registers, then top_k rounds of (lane max -> warp max -> retire one instance).
#include "common.cuh"

#include <mutex>

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

// ---- the same kernel without any allocation ---------------------------------------------------------
// For callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
// that forbid it): the device-side staging of ssp_star_brightness -- image, centres, disc offsets, result --
// lives in ONE file-scope device array of a fixed size instead of the context's growable buffers.  Every GPU
// that loads the library holds these SSP_BRIGHTNESS_STATIC_SCRATCH_BYTES (2.2 MiB) until the process exits.
// In host code the array's NAME is a host-side stand-in, not a device address: the address is asked for
// once per device (cudaGetSymbolAddress with that device current) and only that pointer goes to the kernel.
namespace ssp {

constexpr size_t SBS_IMG_BYTES = (size_t)SSP_BRIGHTNESS_STATIC_MAX_PIXELS;            // uint8 grey image
constexpr size_t SBS_XY_BYTES = (size_t)SSP_BRIGHTNESS_STATIC_MAX_STARS * 8;          // int32 (n, 2)
constexpr size_t SBS_DISC_BYTES = (size_t)32 * SB_MAX_PER_LANE * 8;                   // int32 (n_disc, 2)
constexpr size_t SBS_OUT_BYTES = (size_t)SSP_BRIGHTNESS_STATIC_MAX_STARS * 4;         // int32 (n)
constexpr size_t SBS_XY_OFF = SBS_IMG_BYTES;
constexpr size_t SBS_DISC_OFF = SBS_XY_OFF + SBS_XY_BYTES;
constexpr size_t SBS_OUT_OFF = SBS_DISC_OFF + SBS_DISC_BYTES;
constexpr size_t SBS_TOTAL = SBS_OUT_OFF + SBS_OUT_BYTES;
static_assert(SBS_TOTAL == SSP_BRIGHTNESS_STATIC_SCRATCH_BYTES, "header and scratch layout disagree");
static_assert(SBS_XY_OFF % 256 == 0 && SBS_DISC_OFF % 256 == 0 && SBS_OUT_OFF % 256 == 0, "256-byte aligned parts");

static __device__ __align__(256) unsigned char sbs_scratch[SBS_TOTAL];

constexpr int SBS_MAX_DEVICES = 64;
static std::mutex sbs_mutex;
static unsigned char* sbs_device_ptr[SBS_MAX_DEVICES];   // device address of sbs_scratch, per device; nullptr = not asked yet

// the calling thread's current device must be `device`
static cudaError_t sbs_scratch_address(int device, unsigned char** out) {
  std::lock_guard<std::mutex> lock(sbs_mutex);
  if (sbs_device_ptr[device] == nullptr) {
    void* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, sbs_scratch);
    if (e != cudaSuccess) return e;
    sbs_device_ptr[device] = static_cast<unsigned char*>(p);
  }
  *out = sbs_device_ptr[device];
  return cudaSuccess;
}

}  // namespace ssp

extern "C" int ssp_star_brightness_static(ssp_ctx* ctx, const uint8_t* gray, int h, int w,
                                          const int32_t* center_xy, int64_t n, const int32_t* disc_dxdy,
                                          int n_disc, int top_k, int32_t* brightness, int mem_kind) {
  using namespace ssp;
  if (!ctx) return SSP_ERR_INVALID;
  if (int rc = use_device(ctx)) return rc;
  SSP_REQUIRE(ctx, h > 0 && w > 0 && gray != nullptr, SSP_ERR_INVALID,
              "ssp_star_brightness_static: bad image");
  SSP_REQUIRE(ctx, n >= 0 && (n == 0 || (center_xy != nullptr && brightness != nullptr)),
              SSP_ERR_INVALID, "ssp_star_brightness_static: bad star arrays");
  SSP_REQUIRE(ctx, n_disc > 0 && n_disc <= 32 * SB_MAX_PER_LANE && disc_dxdy != nullptr,
              SSP_ERR_INVALID, "ssp_star_brightness_static: disc must have 1..%d pixels",
              32 * SB_MAX_PER_LANE);
  SSP_REQUIRE(ctx, top_k > 0 && top_k <= n_disc, SSP_ERR_INVALID, "ssp_star_brightness_static: bad top_k");
  SSP_REQUIRE(ctx, (long long)h * w <= (long long)SSP_BRIGHTNESS_STATIC_MAX_PIXELS, SSP_ERR_CAPACITY,
              "ssp_star_brightness_static: image of %d x %d pixels exceeds the fixed scratch (%d pixels); "
              "use ssp_star_brightness",
              h, w, SSP_BRIGHTNESS_STATIC_MAX_PIXELS);
  SSP_REQUIRE(ctx, n <= (int64_t)SSP_BRIGHTNESS_STATIC_MAX_STARS, SSP_ERR_CAPACITY,
              "ssp_star_brightness_static: %lld stars exceed the fixed scratch (%d stars); use "
              "ssp_star_brightness or split the call",
              (long long)n, SSP_BRIGHTNESS_STATIC_MAX_STARS);
  SSP_REQUIRE(ctx, ctx->device >= 0 && ctx->device < SBS_MAX_DEVICES, SSP_ERR_INVALID,
              "ssp_star_brightness_static: device index %d beyond %d", ctx->device, SBS_MAX_DEVICES);
  if (n == 0) return SSP_OK;
  unsigned char* base = nullptr;
  SSP_CUDA_TRY(ctx, sbs_scratch_address(ctx->device, &base));
  unsigned char* d_img = base;
  int* d_xy = reinterpret_cast<int*>(base + SBS_XY_OFF);
  int* d_disc = reinterpret_cast<int*>(base + SBS_DISC_OFF);
  int* d_out = reinterpret_cast<int*>(base + SBS_OUT_OFF);
  // Nothing in the scratch has to start at zero: the kernel reads only what the three copies below have
  // just written and writes out[0, n) completely; what earlier calls left beyond that is never read.
  const cudaMemcpyKind in_kind = to_device_kind(mem_kind);
  SSP_CUDA_TRY(ctx, cudaMemcpyAsync(d_img, gray, (size_t)h * w, in_kind, ctx->stream));
  SSP_CUDA_TRY(ctx, cudaMemcpyAsync(d_xy, center_xy, (size_t)n * 8, in_kind, ctx->stream));
  SSP_CUDA_TRY(ctx, cudaMemcpyAsync(d_disc, disc_dxdy, (size_t)n_disc * 8, cudaMemcpyHostToDevice, ctx->stream));
  const long long threads = n * 32;
  star_brightness_kernel<<<(unsigned)((threads + 127) / 128), 128, 0, ctx->stream>>>(
      d_img, h, w, d_xy, n, d_disc, n_disc, top_k, d_out);
  SSP_CUDA_TRY(ctx, cudaGetLastError());
  SSP_CUDA_TRY(ctx, cudaMemcpyAsync(brightness, d_out, (size_t)n * 4, from_device_kind(mem_kind), ctx->stream));
  SSP_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  return SSP_OK;
}
