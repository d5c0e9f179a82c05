// Context lifetime, stream binding and the descriptor bank of libssp_b200.so.
#include "common.cuh"

static std::string g_create_error;

namespace ssp {
int use_device(ssp_ctx* ctx) {
  cudaError_t e = cudaSetDevice(ctx->device);
  if (e != cudaSuccess)
    return fail(ctx, SSP_ERR_CUDA, "cudaSetDevice(%d) failed: %s", ctx->device,
                cudaGetErrorString(e));
  return SSP_OK;
}
}  // namespace ssp

extern "C" {

int ssp_abi_version(void) { return SSP_ABI_VERSION; }

int ssp_create(ssp_ctx** out, int device) {
  if (!out) {
    g_create_error = "ssp_create: ctx pointer is NULL";
    return SSP_ERR_INVALID;
  }
  *out = nullptr;
  int n_dev = 0;
  cudaError_t e = cudaGetDeviceCount(&n_dev);
  if (e != cudaSuccess || n_dev == 0) {
    g_create_error = std::string("ssp_create: no usable CUDA device (") +
                     (e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0") +
                     "); this engine has no CPU fallback";
    return SSP_ERR_CUDA;
  }
  if (device < 0 || device >= n_dev) {
    g_create_error = "ssp_create: device index out of range";
    return SSP_ERR_INVALID;
  }
  ssp_ctx* ctx = new ssp_ctx();
  ctx->device = device;
  cudaDeviceProp prop;
  if ((e = cudaSetDevice(device)) != cudaSuccess ||
      (e = cudaGetDeviceProperties(&prop, device)) != cudaSuccess) {
    g_create_error = std::string("ssp_create: ") + cudaGetErrorString(e);
    delete ctx;
    return SSP_ERR_CUDA;
  }
  if (prop.major != 10) {
    char buf[160];
    snprintf(buf, sizeof(buf),
             "ssp_create: device %d is sm_%d%d; this library is built for sm_100a (B200) only",
             device, prop.major, prop.minor);
    g_create_error = buf;
    delete ctx;
    return SSP_ERR_CUDA;
  }
  ctx->sm_count = prop.multiProcessorCount;
  int khz = 0;
  cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, device);
  ctx->sm_clock_khz = khz;
  ctx->hbm_bytes = (long long)prop.totalGlobalMem;
  if ((e = cudaStreamCreateWithFlags(&ctx->own_stream, cudaStreamNonBlocking)) != cudaSuccess ||
      (e = cudaEventCreate(&ctx->ev0)) != cudaSuccess ||
      (e = cudaEventCreate(&ctx->ev1)) != cudaSuccess) {
    g_create_error = std::string("ssp_create: ") + cudaGetErrorString(e);
    delete ctx;
    return SSP_ERR_CUDA;
  }
  ctx->stream = ctx->own_stream;
  *out = ctx;
  return SSP_OK;
}

int ssp_destroy(ssp_ctx* ctx) {
  if (!ctx) return SSP_OK;
  cudaSetDevice(ctx->device);
  cudaStreamSynchronize(ctx->stream);
  ssp::DevBuf* bufs[] = {&ctx->desc_owned, &ctx->h_work,      &ctx->h_pairs,     &ctx->h_best2,
                         &ctx->h_counts,   &ctx->h_offsets,   &ctx->h_mq,        &ctx->h_mt,
                         &ctx->h_md,       &ctx->h_expanded,  &ctx->h_popc, &ctx->h_tc_work,  &ctx->h_tc_err,  &ctx->h_cc_aux, &ctx->h_cc_mark, &ctx->h_cc_slots, &ctx->h_cc_cls, &ctx->h_col_key, &ctx->h_col_thr,
                         &ctx->s_xy,        &ctx->s_bright,
                         &ctx->s_rays,     &ctx->p_cand,      &ctx->p_status,    &ctx->p_scan,
                         &ctx->p_tmp_idx,  &ctx->p_tmp_angle, &ctx->p_tmp_side,  &ctx->p_tmp_bright,
                         &ctx->p_tmp_feat, &ctx->p_pairs,     &ctx->p_argmin,    &ctx->p_mindist,
                         &ctx->p_counts,   &ctx->p_offsets,   &ctx->p_m1,        &ctx->p_m2,
                         &ctx->p_md,       &ctx->p_flops,     &ctx->b_img,       &ctx->b_xy,        &ctx->b_disc,
                         &ctx->b_out,      &ctx->c_mask,      &ctx->c_work,      &ctx->c_blocks,    &ctx->c_cand,
                         &ctx->r_pairs,    &ctx->r_kp,        &ctx->r_out,
                         &ctx->t_bank, &ctx->t_na2, &ctx->t_f, &ctx->t_imgs, &ctx->t_work, &ctx->t_na2ptr, &ctx->t_pool, &ctx->t_sv, &ctx->t_keep,
                         &ctx->t_counts, &ctx->t_rowkey, &ctx->t_flags};
  for (ssp::DevBuf* b : bufs) b->release();
  for (ssp::PolyImage& pi : ctx->poly) {
    pi.feat.release();
    pi.star_idx.release();
    pi.angle.release();
    pi.side.release();
    pi.bright.release();
  }
  if (ctx->p_slots) cudaFreeHost(ctx->p_slots);
  if (ctx->ev0) cudaEventDestroy(ctx->ev0);
  if (ctx->ev1) cudaEventDestroy(ctx->ev1);
  if (ctx->own_stream) cudaStreamDestroy(ctx->own_stream);
  delete ctx;
  return SSP_OK;
}

const char* ssp_last_error(const ssp_ctx* ctx) {
  return ctx ? ctx->err.c_str() : g_create_error.c_str();
}

int ssp_set_stream(ssp_ctx* ctx, void* cuda_stream) {
  if (!ctx) return SSP_ERR_INVALID;
  ctx->stream = cuda_stream ? reinterpret_cast<cudaStream_t>(cuda_stream) : ctx->own_stream;
  return SSP_OK;
}

int ssp_synchronize(ssp_ctx* ctx) {
  if (!ctx) return SSP_ERR_INVALID;
  if (int rc = ssp::use_device(ctx)) return rc;
  SSP_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  return SSP_OK;
}

int ssp_host_alloc(ssp_ctx* ctx, int64_t bytes, void** host_ptr) {
  if (!ctx) return SSP_ERR_INVALID;
  SSP_REQUIRE(ctx, host_ptr != nullptr && bytes >= 0, SSP_ERR_INVALID, "ssp_host_alloc: bad arguments");
  if (int rc = ssp::use_device(ctx)) return rc;
  *host_ptr = nullptr;
  SSP_CUDA_TRY(ctx, cudaHostAlloc(host_ptr, (size_t)(bytes > 0 ? bytes : 1), cudaHostAllocPortable));
  return SSP_OK;
}

int ssp_host_free(ssp_ctx* ctx, void* host_ptr) {
  if (!ctx) return SSP_ERR_INVALID;
  if (!host_ptr) return SSP_OK;
  if (int rc = ssp::use_device(ctx)) return rc;
  SSP_CUDA_TRY(ctx, cudaFreeHost(host_ptr));
  return SSP_OK;
}

int ssp_copy_to_host(ssp_ctx* ctx, void* host_dst, const void* device_src, int64_t bytes) {
  if (!ctx) return SSP_ERR_INVALID;
  SSP_REQUIRE(ctx, bytes >= 0 && (bytes == 0 || (host_dst != nullptr && device_src != nullptr)), SSP_ERR_INVALID,
              "ssp_copy_to_host: bad arguments");
  if (int rc = ssp::use_device(ctx)) return rc;
  if (bytes > 0)
    SSP_CUDA_TRY(ctx, cudaMemcpyAsync(host_dst, device_src, (size_t)bytes, cudaMemcpyDeviceToHost, ctx->stream));
  SSP_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  return SSP_OK;
}

int ssp_device_info(ssp_ctx* ctx, int* sm_count, int* sm_clock_khz, int64_t* hbm_bytes) {
  if (!ctx) return SSP_ERR_INVALID;
  if (sm_count) *sm_count = ctx->sm_count;
  if (sm_clock_khz) *sm_clock_khz = ctx->sm_clock_khz;
  if (hbm_bytes) *hbm_bytes = ctx->hbm_bytes;
  return SSP_OK;
}

static int check_offsets(ssp_ctx* ctx, const int64_t* off, int n_images, const char* what,
                         long long max_rows) {
  SSP_REQUIRE(ctx, off != nullptr, SSP_ERR_INVALID, "%s: offsets pointer is NULL", what);
  SSP_REQUIRE(ctx, n_images >= 0, SSP_ERR_INVALID, "%s: n_images < 0", what);
  SSP_REQUIRE(ctx, off[0] == 0, SSP_ERR_INVALID, "%s: offsets[0] must be 0", what);
  for (int i = 0; i < n_images; ++i) {
    SSP_REQUIRE(ctx, off[i + 1] >= off[i], SSP_ERR_INVALID, "%s: offsets must be non-decreasing",
                what);
    SSP_REQUIRE(ctx, off[i + 1] - off[i] < max_rows, SSP_ERR_INVALID,
                "%s: image %d has %lld rows, limit is %lld", what, i,
                (long long)(off[i + 1] - off[i]), max_rows);
  }
  return SSP_OK;
}

int ssp_upload_descriptor_bank(ssp_ctx* ctx, const uint8_t* desc, const int64_t* row_offsets,
                               int n_images, int mem_kind) {
  if (!ctx) return SSP_ERR_INVALID;
  if (int rc = ssp::use_device(ctx)) return rc;
  if (int rc = check_offsets(ctx, row_offsets, n_images, "ssp_upload_descriptor_bank", 1LL << 22))
    return rc;
  const long long rows = row_offsets[n_images];
  SSP_REQUIRE(ctx, rows == 0 || desc != nullptr, SSP_ERR_INVALID,
              "ssp_upload_descriptor_bank: desc is NULL");
  SSP_CUDA_TRY(ctx, ctx->desc_owned.ensure((size_t)rows * 32 + 32));
  if (rows > 0)
    SSP_CUDA_TRY(ctx, cudaMemcpyAsync(ctx->desc_owned.p, desc, (size_t)rows * 32,
                                      ssp::to_device_kind(mem_kind), ctx->stream));
  ctx->d_desc = ctx->desc_owned.as<uint8_t>();
  ctx->row_off.assign(row_offsets, row_offsets + n_images + 1);
  ctx->h_n_matches = -1;
  ctx->h_expanded_valid = false;
  return SSP_OK;
}

int ssp_bind_descriptor_bank(ssp_ctx* ctx, const uint8_t* d_desc, const int64_t* row_offsets,
                             int n_images) {
  if (!ctx) return SSP_ERR_INVALID;
  if (int rc = ssp::use_device(ctx)) return rc;
  if (int rc = check_offsets(ctx, row_offsets, n_images, "ssp_bind_descriptor_bank", 1LL << 22))
    return rc;
  SSP_REQUIRE(ctx, row_offsets[n_images] == 0 || d_desc != nullptr, SSP_ERR_INVALID,
              "ssp_bind_descriptor_bank: d_desc is NULL");
  SSP_REQUIRE(ctx, (reinterpret_cast<uintptr_t>(d_desc) & 15) == 0, SSP_ERR_INVALID,
              "ssp_bind_descriptor_bank: buffer must be 16-byte aligned for TMA bulk copies");
  ctx->d_desc = d_desc;
  ctx->row_off.assign(row_offsets, row_offsets + n_images + 1);
  ctx->h_n_matches = -1;
  ctx->h_expanded_valid = false;
  return SSP_OK;
}

}  // extern "C"
