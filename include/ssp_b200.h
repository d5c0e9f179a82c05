/*
 * ssp_b200.h — C ABI of the B200 pair-matching engine (libssp_b200.so).
 *
 * This is the drop-in boundary for the pairwise image-matching stage of
 * joachim-broser/opencv-starry-sky-panorama-stitcher.  Every entry point names the reference
 * interface it stands in for (paths relative to the reference checkout).  The reference is
 * pure Python, so the binding a maintainer adds is a ctypes stub (INTEGRATION.md shows it).
 *
 * Conventions
 *   - plain C: pointers, sizes, ints.  No torch / cv2 / C++ types cross this boundary.
 *   - every function returns SSP_OK (0) or a negative SSP_ERR_* code; the message is
 *     available from ssp_last_error().  No exceptions cross the ABI.
 *   - `mem_kind` says whether a data pointer is host memory (SSP_MEM_HOST) or memory of the
 *     context's CUDA device (SSP_MEM_DEVICE, e.g. torch.Tensor.data_ptr()).  Index/offset
 *     arrays documented as "host" are always host memory.
 *   - the caller owns every buffer it passes; inputs are read-only and copied (or, for
 *     ssp_bind_descriptor_bank, referenced until the next upload/bind/destroy).
 *   - a context is bound to one CUDA device and is not re-entrant: one host thread per
 *     context (the reference is single-threaded, custom_bruteforce_matcher.py:615-620).
 *   - there is no CPU fallback: if no CUDA device is usable ssp_create fails.
 */
#ifndef SSP_B200_H
#define SSP_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SSP_ABI_VERSION 1

/* status codes */
#define SSP_OK 0
#define SSP_ERR_INVALID (-1)  /* bad argument (null pointer, out-of-range index, misaligned buffer) */
#define SSP_ERR_CUDA (-2)     /* a CUDA runtime call or kernel failed */
#define SSP_ERR_STATE (-3)    /* call order violated (e.g. match before bank upload) */
#define SSP_ERR_CAPACITY (-4) /* caller buffer too small */
#define SSP_ERR_DOMAIN (-5)   /* a polygon hit the reference's math.acos domain error (cbpm:134) */

/* mem_kind */
#define SSP_MEM_HOST 0
#define SSP_MEM_DEVICE 1

/* Hamming match mode */
#define SSP_MODE_KNN2_RATIO 0 /* custom_bruteforce_matcher.py:160-172  knnMatch(k=2) + Lowe ratio  */
#define SSP_MODE_CROSSCHECK 1 /* custom_bruteforce_matcher.py:193-199  BFMatcher(crossCheck=True)   */

/* Hamming kernel variant */
#define SSP_HAMMING_AUTO 0
#define SSP_HAMMING_POPC 1     /* LOP3 + POPC CUDA-core kernel, TMA-bulk staged train tiles         */
#define SSP_HAMMING_POPC_CSA 2 /* same, carry-save-adder compression before POPC (4 POPC/compare)   */
#define SSP_HAMMING_TENSOR 3   /* tcgen05 kind::i8 contraction on int8-expanded descriptors (-+64 query x
                                  0/1 train), accumulators in TMEM; what SSP_HAMMING_AUTO selects   */

/* polygon compare options */
#define SSP_SHIFT_CANONICAL 0       /* parity mode: reference compares shift 0 only (cbpm:992-1016) */
#define SSP_SHIFT_MIN_OVER_SHIFTS 1 /* non-parity: minimum over the 5 cyclic vertex shifts          */
#define SSP_PRECISION_FP64 0        /* parity mode: bit-exact numpy float64 evaluation order         */
#define SSP_PRECISION_FP32 1        /* non-parity fast mode                                          */

/* per-candidate status written by ssp_polygon_measure */
#define SSP_POLY_VALID 0
#define SSP_POLY_SHORT_SIDE 1 /* some pixel distance < min_side_px (cbpm:515)        */
#define SSP_POLY_NOT_CONVEX 2 /* not all 5 stars are hull vertices (cbpm:520-526)    */
#define SSP_POLY_DOMAIN 3     /* |acos argument| > 1 (cbpm:134 raises ValueError)    */
#define SSP_POLY_BAD_INDEX 4  /* a star index is outside the image's star table        */

typedef struct ssp_ctx ssp_ctx;

/* ---- lifetime ------------------------------------------------------------------------ */

/* Creates an engine context on CUDA device `device`.  One per CustomBruteForceMatcher instance
 * (it replaces the per-matcher state custom_bruteforce_matcher.py:57-59 keeps). */
int ssp_create(ssp_ctx** ctx, int device);
int ssp_destroy(ssp_ctx* ctx);
/* Last error message of `ctx` (or of the last failed ssp_create when ctx == NULL). */
const char* ssp_last_error(const ssp_ctx* ctx);
int ssp_abi_version(void);
/* Run all subsequent work on `cuda_stream` (a cudaStream_t, e.g. torch's current stream
 * handle).  NULL selects the context's own stream. */
int ssp_set_stream(ssp_ctx* ctx, void* cuda_stream);
int ssp_synchronize(ssp_ctx* ctx);
/* Device facts used by the roofline report: SM count and SM clock in kHz (max). */
int ssp_device_info(ssp_ctx* ctx, int* sm_count, int* sm_clock_khz, int64_t* hbm_bytes);
/* Page-locked host memory for caller-allocated input / output buffers (descriptor banks going
 * in, survivor arrays coming out): copies to and from it run at full PCIe / C2C rate and truly
 * asynchronously, where pageable numpy memory is staged by the driver.  The reference has no
 * counterpart (its arrays never leave the host); optional, any host pointer is accepted by the
 * MEM_HOST paths.  Free with ssp_host_free. */
int ssp_host_alloc(ssp_ctx* ctx, int64_t bytes, void** host_ptr);
int ssp_host_free(ssp_ctx* ctx, void* host_ptr);
/* Copies `bytes` from an engine-owned device table (an address an entry point handed out, e.g.
 * ssp_polygon_candidates) to host memory, ordered after the work queued on the context's stream. */
int ssp_copy_to_host(ssp_ctx* ctx, void* host_dst, const void* device_src, int64_t bytes);

/* ---- descriptor bank ------------------------------------------------------------------
 * Replaces the per-pair `ft.descriptors.get()[0:n_orb, :]` host slices handed to
 * cv.BFMatcher at custom_bruteforce_matcher.py:197-198.  `desc` holds the ORB rows of all
 * images back to back, 32 bytes per row (uint8 (N,32) C-contiguous); image i owns rows
 * [row_offsets[i], row_offsets[i+1]).  `row_offsets` is host memory, n_images+1 entries.
 * Rows per image must be < 2^22 (OpenCV itself asserts < 2^18 per train set). */
int ssp_upload_descriptor_bank(ssp_ctx* ctx, const uint8_t* desc, const int64_t* row_offsets,
                               int n_images, int mem_kind);
/* Zero-copy variant: use the caller's device buffer in place (must be 16-byte aligned and stay
 * alive and unmodified until the next upload/bind or ssp_destroy). */
int ssp_bind_descriptor_bank(ssp_ctx* ctx, const uint8_t* d_desc, const int64_t* row_offsets,
                             int n_images);

/* ---- kernel 1: Hamming all-pairs ------------------------------------------------------
 * Replaces cv.BFMatcher(cv.NORM_HAMMING[, crossCheck=True]).match / .knnMatch(k=2) + ratio
 * test for a whole list of image pairs in one call (the pair loop
 * custom_bruteforce_matcher.py:615-620).  `pairs` is host memory, n_pairs x 2 int32
 * (query image, train image).
 *
 *   SSP_MODE_KNN2_RATIO : per query row the two nearest train rows in the order of a stable
 *       ascending sort (lowest train index first among equal distances); a query survives iff
 *       ratio_den * d1 < ratio_num * d2 (0.75 -> 3, 4; exact integer form of :171).
 *       Train sets with fewer than 2 rows produce no survivors (the reference's `for m, n in`
 *       would raise; the Python layer raises ValueError for that case).
 *   SSP_MODE_CROSSCHECK : (q, t) survives iff t = argmin_t D[q,.] and q = argmin_q D[.,t],
 *       both with the lowest index on ties (OpenCV batchDistance crosscheck semantics).
 *       The tensor variant searches the reverse direction only for the train rows some query
 *       row claims (same result; environment SSP_B200_CROSSCHECK_DENSE=1 searches all of them).
 *
 * Survivors of pair p are reported in ascending query index (OpenCV's order) at positions
 * [match_offsets[p], match_offsets[p+1]) of the arrays ssp_hamming_fetch_matches returns.
 * `match_offsets` is host memory with n_pairs+1 entries.  The call synchronises the stream.
 * A train set (in cross-check mode either image) must have fewer than 2^18 rows, the limit
 * cv::BFMatcher itself asserts (IMGIDX_SHIFT); more is SSP_ERR_INVALID. */
int ssp_hamming_match_pairs(ssp_ctx* ctx, const int32_t* pairs, int n_pairs, int mode,
                            int ratio_num, int ratio_den, int variant, int64_t* match_offsets);
/* The same with the Lowe ratio as the float64 the reference multiplies with
 * (custom_bruteforce_matcher.py:164 `m.distance < 0.75 * n.distance`): the device evaluates
 * (double)d1 < ratio * (double)d2 with one correctly rounded multiplication, i.e. exactly Python's
 * comparison for ANY ratio.  ssp_hamming_match_pairs(num, den) is this call with
 * ratio = (double)num / (double)den. */
int ssp_hamming_match_pairs_f64(ssp_ctx* ctx, const int32_t* pairs, int n_pairs, int mode,
                                double ratio, int variant, int64_t* match_offsets);
/* Copies the survivors of the last ssp_hamming_match_pairs call: queryIdx, trainIdx and the
 * integer Hamming distance (DMatch.distance is this value as float32).  `capacity` is the
 * number of entries each array can hold (>= match_offsets[n_pairs]). */
int ssp_hamming_fetch_matches(ssp_ctx* ctx, int32_t* query_idx, int32_t* train_idx,
                              int32_t* distance, int64_t capacity, int mem_kind);
/* Optional: the raw per-query nearest neighbours of the last call, one int32[4] record
 * {idx1, dist1, idx2, dist2} per query row, pairs back to back in call order (for
 * SSP_MODE_CROSSCHECK the forward direction only; idx2/dist2 are -1 there and whenever the
 * train set has a single row).  `capacity_rows` >= sum of query rows over the pairs. */
int ssp_hamming_fetch_best2(ssp_ctx* ctx, int32_t* best2, int64_t capacity_rows, int mem_kind);
/* Number of kernel launches and device time (ms, CUDA events on the context's stream) of the
 * dominant Hamming kernel during the last ssp_hamming_match_pairs call. */
int ssp_hamming_last_stats(ssp_ctx* ctx, int* n_launches, float* main_kernel_ms,
                           int64_t* comparisons);

/* ---- kernel 2: star -> sphere projection and polygon measure --------------------------
 * Star table of all images (the `keypoints[n_orb:]` star keypoints of
 * custom_bf_polygon_matcher.py:710 plus their get_star_brightness values,
 * image_processors.py:222-254).  xy: float64 (n_stars, 2) pixel coordinates; brightness:
 * float64 (n_stars); image i owns stars [star_offsets[i], star_offsets[i+1]) (host array).
 * img_w/img_h/focal_px are the pinhole intrinsics of angle_distance (cbpm:86-108): each star
 * becomes the unit ray of (x - int(img_w/2), y - int(img_h/2), focal_px).  Replaces the whole
 * star table; polygon banks measured earlier are kept (they own copies of their measures). */
int ssp_upload_star_bank(ssp_ctx* ctx, const double* xy, const double* brightness,
                         const int64_t* star_offsets, int n_images, int img_w, int img_h,
                         double focal_px, int mem_kind);
/* Reads back the unit rays (n_stars x 3 float64) computed at upload. */
int ssp_fetch_star_rays(ssp_ctx* ctx, double* rays, int64_t capacity_stars, int mem_kind);

/* Measures the candidate 5-star polygons of one image: the per-candidate body of
 * get_polygons_characteristics (cbpm:500-600) and the normalisation of cbpm:722-740.
 * `candidates`: int32 (n_candidates, 5) star indices local to `image`.  Valid polygons are
 * stored, in candidate order, in the context's polygon bank for that image (replacing
 * polygon_data_store[img_id], cbpm:744-748).  `status` (optional, host, n_candidates bytes)
 * receives SSP_POLY_* per candidate; `n_valid` receives the number of stored polygons.
 * Returns SSP_ERR_DOMAIN (after storing everything else) if any candidate hit the acos domain
 * error on which the reference raises.
 * With `status` == NULL and `n_valid` == NULL the call does not wait for the device (a batch of
 * images is then measured back to back); the count -- and a domain error, if any -- is delivered
 * by the next ssp_polygon_count / _fetch / _match_pairs call. */
int ssp_polygon_measure(ssp_ctx* ctx, int image, const int32_t* candidates, int64_t n_candidates,
                        double min_side_px, int mem_kind, uint8_t* status, int64_t* n_valid);
/* The same for several images in ONE launch pair (measure + compaction): images[k] owns the
 * candidates [cand_offsets[k], cand_offsets[k+1]) of the concatenated int32 (n, 5) table.  Always the
 * no-wait form: counts and domain errors are delivered by the next ssp_polygon_count / _fetch /
 * _match_pairs call.  A step of the star path measures every image of the panorama this way
 * (the reference: once per image, cached in polygon_data_store, cbpm:744-748). */
int ssp_polygon_measure_batch(ssp_ctx* ctx, const int32_t* images, int n_images, const int32_t* candidates,
                              const int64_t* cand_offsets, double min_side_px, int mem_kind);
/* Reads back the stored polygons of `image`, any pointer may be NULL.  All arrays are
 * (P, 5) row-major in the canonical vertex order (hull order, orientation of cbpm:537-543,
 * rotated so the smallest vertex angle is element 0, cbpm:578-591):
 *   star_idx int32, angle / side float64 in radians (un-normalised), brightness float64
 *   (un-normalised), feat float64 (P, 15) = [angle/pi, side/max side, brightness/max
 *   brightness] exactly as the reference stores them after cbpm:722-740. */
int ssp_polygon_fetch(ssp_ctx* ctx, int image, int32_t* star_idx, double* angle, double* side,
                      double* brightness, double* feat, int64_t capacity, int mem_kind);
int ssp_polygon_count(ssp_ctx* ctx, int image, int64_t* n_polygons);
/* Loads a polygon bank directly (normalised 15-dim rows + star indices, which may be NULL);
 * used for synthetic banks and for resuming from a pickled polygon_data_store. */
int ssp_polygon_upload_features(ssp_ctx* ctx, int image, const double* feat,
                                const int32_t* star_idx, int64_t n_polygons, int mem_kind);

/* ---- kernel 3: polygon versus polygon -------------------------------------------------
 * Replaces the all-pairs distance block cbpm:992-1034, the row argmin cbpm:1063-1067 and
 * the threshold cbpm:1145-1148 for a list of image pairs.  For each polygon r of the first
 * image: c* = first argmin_c dist[r,c]; (r, c*, dist) is reported iff dist < threshold
 * (reference: 0.1566), in ascending r.  SSP_PRECISION_FP64 + SSP_SHIFT_CANONICAL reproduces
 * numpy's float64 result bit for bit (same summation order, no FMA contraction). */
int ssp_polygon_match_pairs(ssp_ctx* ctx, const int32_t* pairs, int n_pairs, double threshold,
                            int shift_mode, int precision, int64_t* match_offsets);
int ssp_polygon_fetch_matches(ssp_ctx* ctx, int32_t* poly1, int32_t* poly2, double* dist,
                              int64_t capacity, int mem_kind);
/* Optional: per polygon of the first image of every pair, the argmin column and minimum
 * distance before thresholding (pairs back to back). */
int ssp_polygon_fetch_rowmin(ssp_ctx* ctx, int32_t* argmin, double* mindist,
                             int64_t capacity_rows, int mem_kind);
int ssp_polygon_last_stats(ssp_ctx* ctx, int* n_launches, float* main_kernel_ms,
                           int64_t* evaluations);
/* The parity compare prunes exactly (a column whose partial sum already reaches the row's best is
 * dropped): flops it actually executed in the last call, counted on the device, or -1 when the
 * kernel that ran does not count (fp32 / min-over-shifts / legacy). */
int ssp_polygon_last_executed_flops(ssp_ctx* ctx, int64_t* executed_flops);
/* The same count split by precision: the compacted-pruning kernel's first-group FILTER runs in single
 * precision (conservative: it only decides what the FP64 evaluation looks at), everything that
 * produces a result in FP64.  Both -1 when the kernel that ran does not count. */
int ssp_polygon_last_executed_flops_by_precision(ssp_ctx* ctx, int64_t* fp64_flops, int64_t* fp32_flops);
/* The default parity compare screens every (row, column) cell with a TF32 tensor-core contraction and
 * evaluates in FP64 only the cells that can hold a row's minimum.  Of the last call: whether that form ran
 * to the end (used_tensor), why it was abandoned for the CUDA-core kernels if it was (fallback_flags: 1
 * shared-memory alignment, 2 features beyond the float range, 4 candidate pool overflow, 8 a screen value
 * further from its FP64 value than the bound), the candidates it evaluated in FP64, the largest
 * abs(screen - FP64) / F^2 seen among them and the bound it must stay under (F = largest abs feature). */
int ssp_polygon_last_screen_stats(ssp_ctx* ctx, int* used_tensor, int* fallback_flags, int64_t* candidates,
                                  double* worst_error_coef, double* bound_coef);

/* ---- row N3: batched RANSAC homography (opt-in, NOT bit-compatible with OpenCV) -----------
 * Stands in for the per-pair cv.findHomography(src_pts, dst_pts, cv.RANSAC, 5) of
 * custom_bruteforce_matcher.py:215-246 for ALL pairs of the last ssp_hamming_match_pairs call at
 * once; the point pairs are gathered on the device from that call's survivors (query row ->
 * keypoint of the query image, train row -> keypoint of the train image).
 *   keypoint_xy       float32 (K, 2) pixel coordinates of all images' keypoints back to back (host)
 *   keypoint_offsets  int64 (n_images + 1) (host); image i must list at least as many keypoints as
 *                     it has descriptor rows
 *   pairs / n_pairs   the pair list of the last Hamming call
 *   reproj_threshold  inlier bound in pixels (reference: 5); n_hypotheses 4-point samples per pair
 *   min_matches       pairs with fewer matches get no model (reference: 6, :215): H = NaN, n_inliers = -1
 *   H                 float64 (n_pairs, 9) row-major, h33 = 1: least-squares refit on the inliers of
 *                     the best hypothesis (NaN if no model)
 *   inlier_mask       uint8 per survivor, in the order of ssp_hamming_fetch_matches
 *   n_inliers         int32 per pair
 * Same model as cv::findHomography (4-point hypotheses, reprojection error, best count wins, refit on
 * the inliers, mask of the best hypothesis) with a fixed seed and hypothesis count; OpenCV's RNG,
 * adaptive termination and Levenberg-Marquardt polish are not reproduced. */
int ssp_ransac_homography_pairs(ssp_ctx* ctx, const float* keypoint_xy, const int64_t* keypoint_offsets,
                                int n_images, const int32_t* pairs, int n_pairs, double reproj_threshold,
                                int n_hypotheses, uint32_t seed, int min_matches, double* H,
                                uint8_t* inlier_mask, int64_t mask_capacity, int32_t* n_inliers);

/* ---- row N2: candidate generation ------------------------------------------------------
 * The data-parallel parts of reduce_amount_of_keypoints_and_create_polygon_spanning_tuples_of_keypoints
 * (cbpm:261-443) and reduce_amount_of_keypoints_by_clustering (cbpm:137-259); the Ward clustering
 * (sklearn, cbpm:175-181) stays on the host and its labels are an input.
 *
 * ssp_star_regions: bit r of region_mask[i] = star i of `image` (star bank) lies in region r of
 * cbpm:328-375, regions in the reference's dict order (left_top, left_bot, left_mid, right_top,
 * right_bot, right_mid, top_left, top_right, top_mid, bot_left, bot_right, bot_mid); the strict
 * float64 comparisons of the reference, evaluated in its order. */
int ssp_star_regions(ssp_ctx* ctx, int image, uint16_t* region_mask, int64_t capacity);
/* cbpm:230-245 for several regions at once: region g owns members[region_offsets[g] ..
 * region_offsets[g+1]) (star ids of `image`, in the order the host clustered them) with their
 * cluster labels (0 .. max_labels-1); survivors[g * max_labels + l] = the brightest star of cluster
 * l of region g, the LAST listed among equally bright ones, or -1 for an unused label. */
int ssp_cluster_brightest(ssp_ctx* ctx, int image, const int32_t* members, const int32_t* labels,
                          const int32_t* region_offsets, int n_regions, int max_labels,
                          int32_t* survivors);
/* cbpm:407-413: set(itertools.combinations(region list, 5) for each of the 12 regions).
 * region_members: the regions' star ids back to back IN LIST ORDER (the order decides which tuples
 * are equal), region_counts[12].  *n_tuples = number of distinct tuples, i.e. len() of the reference's
 * set, the number its reduction loop compares with the maximum (five stars listed by two regions in a
 * different relative order count twice); *n_unique = number of distinct star SETS.  If
 * 0 < *n_tuples <= max_candidates and device_candidates != NULL, one tuple per distinct star set is
 * written to an engine-owned int32 (n_unique, 5) table on the device -- region-major, combination
 * order inside a region; a set listed by several regions belongs to the first -- and its address is
 * returned for ssp_polygon_measure(..., SSP_MEM_DEVICE); otherwise *device_candidates = NULL (the
 * host retries with fewer stars per region, as the reference's while loop does). */
int ssp_polygon_candidates(ssp_ctx* ctx, const int32_t* region_members, const int32_t* region_counts,
                           int n_regions, int64_t max_candidates, int64_t* n_tuples, int64_t* n_unique,
                           const int32_t** device_candidates);

/* ---- row N1: star brightness ----------------------------------------------------------
 * get_star_brightness (image_processors.py:222-254) for many stars of one image:
 * sum of the 10 largest gray values inside the filled circle cv.circle(center=(int(x), int(y)),
 * radius=10).  `gray`: uint8 (h, w), the cv.cvtColor(BGR2GRAY) image the reference recomputes
 * for every star (the host converts once per image with OpenCV itself); `center_xy`: int32
 * (n, 2) already truncated like the reference's int(); `disc_dxdy`: int32 (n_disc, 2) pixel
 * offsets of the rasterised disc (the host obtains them from cv.circle once, so the mask is
 * OpenCV's own); result int32 (n). */
int ssp_star_brightness(ssp_ctx* ctx, const uint8_t* gray, int h, int w, const int32_t* center_xy,
                        int64_t n, const int32_t* disc_dxdy, int n_disc, int top_k,
                        int32_t* brightness, int mem_kind);

/* ssp_star_brightness without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid it).  Same arguments, same results; the device-side staging (image,
 * centres, disc offsets, result) lives in one file-scope device array of
 * SSP_BRIGHTNESS_STATIC_SCRATCH_BYTES that every GPU which loads the library holds until the process
 * exits.  The call itself performs no cudaMalloc / cudaFree and calls no allocating library routine.
 * Limits: h * w <= SSP_BRIGHTNESS_STATIC_MAX_PIXELS (2 Mi pixels; the reference's 894 x 1342 work image is
 * 1.2 M), n <= SSP_BRIGHTNESS_STATIC_MAX_STARS; beyond them SSP_ERR_CAPACITY with a message, nothing
 * launched.  ALL calls on one GPU (from any context or thread) share the array: it keeps the last call's
 * contents, nothing in it needs clearing between calls, and two calls that overlap on the GPU race -- run
 * them one after the other (each call returns after its stream has drained). */
#define SSP_BRIGHTNESS_STATIC_MAX_PIXELS (2 * 1024 * 1024)
#define SSP_BRIGHTNESS_STATIC_MAX_STARS 16384
#define SSP_BRIGHTNESS_STATIC_SCRATCH_BYTES (SSP_BRIGHTNESS_STATIC_MAX_PIXELS + SSP_BRIGHTNESS_STATIC_MAX_STARS * 12 + 4096)
int ssp_star_brightness_static(ssp_ctx* ctx, const uint8_t* gray, int h, int w, const int32_t* center_xy,
                               int64_t n, const int32_t* disc_dxdy, int n_disc, int top_k,
                               int32_t* brightness, int mem_kind);

#ifdef __cplusplus
}
#endif
#endif /* SSP_B200_H */
