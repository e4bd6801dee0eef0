/*
 * rct.h -- C ABI of librct.so: B200-native regional colour transfer.
 *
 * Drop-in boundary for the hot path of the reference
 *   Style_3DGS/localized_style_transfer.py  (abbrev. LST), lines 11-168 and 222-238.
 * Every entry point names the reference interface it replaces.  Signatures carry plain
 * pointers and sizes only (no torch / numpy types).  Unless a name ends in `_host`, data
 * pointers are DEVICE pointers and the call is asynchronous on `stream` (a cudaStream_t
 * passed as void*; NULL = the legacy default stream).  No entry point synchronises the
 * device except where stated.  There is no CPU fallback: without a CUDA device every
 * compute entry point returns RCT_ERR_CUDA.
 *
 * Image layout: uint8, HWC, C-contiguous (what np.array(PIL.Image) yields, LST:200-201);
 * batches are NHWC with explicit frame strides in bytes.
 *
 * Return convention: 0 = ok, 1 = empty target region ("Warning: No
 * foreground pixels found", LST:141-143), 2 = empty source region (LST:145-147), <0 = error
 * (rct_last_error() gives the text).  On 1/2 the output is what the reference returns in
 * that case (the unmodified foreground composited with the background).
 */
#ifndef RCT_H_
#define RCT_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RCT_OK 0
#define RCT_EMPTY_TARGET 1
#define RCT_EMPTY_SOURCE 2
#define RCT_ERR_CUDA (-1)
#define RCT_ERR_ARG (-2)
#define RCT_ERR_NOMEM (-3)

/* Region selection modes of the fused pipeline.
 * MASK : `mask` is the reference's background_mask (LST:218), values {0,1}.  Target =
 *        content pixels with mask==0, source = stylised pixels with mask!=0; the output is
 *        LST:238's combined_image.
 * LABEL: `mask` is a label map; target = content, source = stylised, both restricted to
 *        mask==label; only those output pixels are written (one pass per label; the 8-label
 *        configuration of BASELINE.json, which has no counterpart in the reference).
 * ALL  : color_transfer_foreground(content, stylised) itself (LST:128-168); `mask` ignored. */
#define RCT_MODE_MASK 0
#define RCT_MODE_LABEL 1
#define RCT_MODE_ALL 2

#define RCT_FLAG_DEBUG 1u /* keep per-colour rank / matched-value tables for rct_plan_debug_* */

typedef struct rct_plan rct_plan; /* opaque; one per (device, geometry); not thread-safe */

int rct_version(void);
const char *rct_last_error(void); /* thread-local text of the last error on this thread */

/* a-0: the four Reinhard matrices RGB_TO_LMS, LMS_TO_LAB, LAB_TO_LMS, LMS_TO_RGB (LST:12-19),
 * 4 x 9 doubles, row-major.  The host computes them with numpy exactly as LST:12-19 does
 * (the inverses carry LAPACK round-off) and installs them here; defaults are the values of
 * numpy 2.3.5 / OpenBLAS 0.3.30.  Takes effect for plans created afterwards. */
int rct_set_matrices(const double *m36);
int rct_get_matrices(double *m36);

/* a-1: Pillow Image.resize(..., Image.NEAREST) source-index table (LST:223-230): host
 * function, writes out_size int32 indices (accumulated-double recurrence). */
int rct_resample_table(int in_size, int out_size, int32_t *tab_host);

/* ---- plans ----------------------------------------------------------------------------
 * A plan fixes the geometry: content/mask/output are H x W, the stylised image is sh x sw
 * (resampled to H x W on the fly with the a-1 tables when different, never materialised),
 * and `slots` frames are processed per launch group.  `workspace` is device memory of at
 * least rct_plan_workspace_bytes() bytes owned by the caller (e.g. a torch tensor), or
 * NULL to let the library cudaMalloc it.  Creation synchronises the device once. */
size_t rct_plan_workspace_bytes(int H, int W, int slots, unsigned flags);
int rct_plan_create(int device, int H, int W, int sh, int sw, int slots, unsigned flags,
                    void *workspace, size_t workspace_bytes, rct_plan **plan);
int rct_plan_destroy(rct_plan *plan);
/* tuning knobs; RCT_OPT_AGGREGATE (0/1): warp-aggregate equal colours before the histogram
 * atomic (pays off on flat / low-colour images, costs a MATCH per pixel otherwise) */
#define RCT_OPT_AGGREGATE 0
/* RCT_OPT_TIMING (0/1/2): 1 = bracket the phases of every launch group of rct_transfer_async /
 * rct_transfer_host with CUDA events on the launching stream; read with rct_plan_timing() */
#define RCT_OPT_TIMING 1
/* RCT_OPT_PART_CAP (1..4096): largest sort part handled in shared memory; lowering it forces the
 * global-memory fallback of the exact sort (tests only) */
#define RCT_OPT_PART_CAP 2
/* RCT_OPT_FRONTEND: how pixels become the per-region (colour, count) lists.  1 (default): colour
 * partition + shared-memory histograms (streaming passes only); 0: the two 2^24-bin colour-cube
 * histograms in device memory (what the row-sharded path all-reduces).  Same lists either way. */
#define RCT_OPT_FRONTEND 3
/* RCT_OPT_APPLY: the apply pass when the call is MASK mode, not resampled and 16-byte aligned.
 * 1 (default): one bulk-async (TMA 1-D, cp.async.bulk + mbarrier) staged pipeline per warp;
 * 2: the same pipeline fed by cp.async (LDGSTS); 3: one CTA-wide bulk-async ring; 4: 512-pixel tiles;
 * 5: five CTAs per SM; 6: three stages; 7: output through shared memory + one bulk store per tile
 * (A/B variants, all measured slower or equal: profiles/r01_summary_v10.md); 0: always the register-staged
 * kernel (which also serves every other geometry).
 * 8: as 1, and when the stylised image is resampled (W >= 256, 16-byte aligned frames) the NEAREST
 * resample is staged per tile by bulk copies of the source row spans (k_apply_warp_rs); measured slower
 * than the register-staged gather (the source image is L2-resident), hence opt-in. */
#define RCT_OPT_APPLY 4
/* RCT_OPT_MATCH: traversal of a sorted target part by the CDF-match kernel.  1 (default): lanes
 * interleaved (coalesced, every search starts from an interpolated position); 0: one contiguous
 * run per thread with hinted searches.  Same results. */
#define RCT_OPT_MATCH 5
/* RCT_OPT_STAGE_SETS (2..4, default 4): device staging sets of rct_transfer_host; group k is copied
   in while group k-1 computes and group k-2 is copied out. */
#define RCT_OPT_STAGE_SETS 6
/* RCT_OPT_LUT: layout of the per-frame colour -> output table.  0 (default): uint32 RGB per colour (64 MiB);
 * 1: uint16 run index per colour (32 MiB, L2-resident during the apply pass) + palette of the run colours --
 * the apply pass gets 1.5x faster on noisy images, the three extra solve kernels cost more than that
 * (measured, profiles/r01_summary_v10.md), so it stays opt-in. */
#define RCT_OPT_LUT 7
/* RCT_OPT_KEYTAB (default 1): the key table in front of the colour -> RGB table (csrc/rct_keytab.cuh).  The output
 * colour is a function of the colour's projection key alone; an fp32 key with a proven error bound picks the run of
 * equal output colour from a ~300 KB L1/L2-resident cell table whenever no run boundary lies within the bound, and only
 * the remaining ~10 % of the target pixels gather from the 64 MiB table.  Bit-identical results (every GPU parity test
 * runs with it; rct_plan_debug_keytab re-checks every distinct target colour); 0 = colour table only (A/B runs). */
#define RCT_OPT_KEYTAB 8
int rct_plan_set_option(rct_plan *plan, int option, int value);
/* accumulated device time per phase since the last reset, in ms: [0] workspace memsets,
 * [1] k_hist (pixels -> colour histograms), [2] solve (compact .. match -> LUT), [3] k_apply
 * (pixels -> output); *groups = launch groups accumulated.  Synchronises the recorded events. */
int rct_plan_timing(rct_plan *plan, double *ms4, long long *groups, int reset);
/* RCT_OPT_TIMING = 2: one CUDA event after EVERY kernel of rct_transfer_async (on the launching
 * stream); rct_plan_kernel_timing() returns the accumulated device time per kernel id since the
 * last reset, ms[RCT_TK_COUNT], and the launch groups accumulated.  A group's first interval
 * (id RCT_TK_BEGIN) is not counted; memsets are their own id. */
enum {
    RCT_TK_BEGIN = 0, RCT_TK_MEMSET, RCT_TK_FRONT_COUNT, RCT_TK_FRONT_SCAN, RCT_TK_FRONT_PART, RCT_TK_FRONT_HIST,
    RCT_TK_HIST, RCT_TK_COMPACT, RCT_TK_LAB_MOMENTS, RCT_TK_PCA, RCT_TK_KEYS, RCT_TK_PLAN_FINE, RCT_TK_FINE_HIST,
    RCT_TK_PLAN_PARTS, RCT_TK_SCATTER, RCT_TK_PART_SORT, RCT_TK_PART_SCAN, RCT_TK_MATCH, RCT_TK_APPLY, RCT_TK_KEYTAB,
    RCT_TK_COUNT
};
int rct_plan_kernel_timing(rct_plan *plan, double *ms, long long *groups, int reset);
const char *rct_kernel_timing_name(int id);

/* ---- the fused pipeline: LST:222-238 + color_transfer_foreground (LST:128-168) ----------
 * n_frames frames (any number; processed `slots` at a time), frame f at base + f*stride.
 * stride 0 for `stylised`/`mask` broadcasts one image to all frames.  `mask` may be NULL in
 * RCT_MODE_ALL.  Per-frame status codes are left in plan memory; rct_plan_status() fetches
 * them (synchronises `stream`). */
int rct_transfer_async(rct_plan *plan, const uint8_t *content, size_t content_stride,
                       const uint8_t *stylised, size_t stylised_stride, const uint8_t *mask,
                       size_t mask_stride, uint8_t *out, size_t out_stride, int n_frames,
                       int mode, int label, void *stream);
int rct_plan_status(rct_plan *plan, int *status_host, int n_frames, void *stream);

/* BASELINE.json configs[4] (no counterpart in the reference, which knows one foreground / background pair): for every
 * label l < n_labels of a uint8 label map, the content pixels with label l are matched to the stylised pixels with label
 * l (LABEL mode of rct_transfer_async), all labels of slots / n_labels frames per launch group side by side.  The plan
 * needs slots >= n_labels.  `out` must be pre-filled with the values untouched pixels keep (a copy of content).
 * Status per (frame, label) lands in the plan's status array (rct_plan_status, n = n_frames * n_labels). */
int rct_transfer_labels_async(rct_plan *plan, const uint8_t *content, size_t content_stride, const uint8_t *stylised,
                              size_t stylised_stride, const uint8_t *labels, size_t labels_stride, uint8_t *out,
                              size_t out_stride, int n_frames, int n_labels, void *stream);
/* Same through HOST buffers (the reference's calling convention: numpy arrays in, numpy
 * array out).  Copies in, runs, copies out, synchronises; pinned host memory makes the
 * copies true DMA.  Returns the worst per-frame status. */
int rct_transfer_host(rct_plan *plan, const uint8_t *content_h, const uint8_t *stylised_h,
                      const uint8_t *mask_h, uint8_t *out_h, int n_frames, int mode, int label,
                      int *status_host /* [n_frames] or NULL */);

/* ---- row-sharded single image: the pipeline in three phases ---------------
 * Each rank owns rows [row0, row0+rows) of the H x W image described by the plan (slot 0).
 * 1. rct_shard_hist_async   : zero the colour-cube histograms and count the local rows;
 * 2. all-reduce (sum, uint32) the buffer rct_plan_hist_ptr()/rct_plan_hist_bytes() over the
 *    ranks -- done by the caller (NCCL through torch.distributed);
 * 3. rct_shard_solve_async  : statistics, PCA, exact CDF matching -> colour LUT (redundant,
 *    identical on every rank);
 * 4. rct_shard_apply_async  : write the local rows of the output.
 * Pointers are to the FULL images (row offsets are applied inside). */
int rct_shard_hist_async(rct_plan *plan, const uint8_t *content, const uint8_t *stylised,
                         const uint8_t *mask, int mode, int label, int row0, int rows,
                         void *stream);
void *rct_plan_hist_ptr(rct_plan *plan);
size_t rct_plan_hist_bytes(rct_plan *plan);
int rct_shard_solve_async(rct_plan *plan, void *stream);
int rct_shard_apply_async(rct_plan *plan, const uint8_t *content, const uint8_t *stylised,
                          const uint8_t *mask, uint8_t *out, int mode, int label, int row0,
                          int rows, void *stream);

/* ---- BINNED mode: B-bin key histograms + CDF build / inverse lookup (BASELINE.json configs[4] "1024-bin
 * histograms"; north_star subsystems 2-3) ---------------------------------------------------------------------------
 * NO COUNTERPART IN THE REFERENCE: match_cdf (LST:99-125) is exact and sort-based, and rct_transfer_async reproduces
 * it exactly.  This mode replaces only that step by classic histogram matching (equal-width bins over the region's
 * key range, in-bin rank lerp, inverse lookup in the source CDF; definition in csrc/rct_binned.cuh and
 * oracle/rct_oracle.c:orc_match_binned); everything else (a-1..a-5, a-7..a-9) is the exact path's arithmetic.  It is an
 * APPROXIMATION (measured: within 1 uint8 LSB of the exact result on all but isolated tail pixels (max 3-4 observed); 0.5-5 % of the pixels differ), bit-exact against its own
 * oracle, never the parity path.  2 <= bins <= 4096.  Same arguments as rct_transfer_async otherwise. */
int rct_transfer_binned_async(rct_plan *plan, const uint8_t *content, size_t content_stride, const uint8_t *stylised,
                              size_t stylised_stride, const uint8_t *mask, size_t mask_stride, uint8_t *out,
                              size_t out_stride, int n_frames, int mode, int label, int bins, void *stream);
/* Row-sharded binned path ("the binned approximate mode ... two tiny all-reduces"): the only cross-GPU
 * state is the exchange buffer rct_plan_xchg_ptr(), which the caller all-reduces between the phases:
 *   phase 0: band -> colour lists -> exact integer moments     then SUM  over int64  [56]       (which = 0)
 *   phase 1: PCA from the global moments -> keys -> key range  then MAX  over double [4]        (which = 1)
 *   phase 2: histograms over the global range                  then SUM  over uint32 [2*bins]   (which = 2)
 *   phase 3: CDFs -> LUT of the band's colours -> the band's rows of `out`
 * Integer sums and float max are order-independent: the result is bit-identical for any number of ranks. */
int rct_shard_binned_phase_async(rct_plan *plan, int phase, const uint8_t *content, const uint8_t *stylised,
                                 const uint8_t *mask, uint8_t *out, int mode, int label, int row0, int rows, int bins,
                                 void *stream);
void *rct_plan_xchg_ptr(rct_plan *plan);
size_t rct_plan_xchg_bytes(int which, int bins);
/* introspection for the parity tests: histograms [bins], key ranges (lo_t hi_t lo_s hi_s) and, for plans created with
 * RCT_FLAG_DEBUG, per entry of the target colour list its colour, CDF index pair b * 65536 + b' and matched value
 * (arrays of rct_frame_info.d_target entries; any pointer may be NULL) */
int rct_plan_debug_binned(rct_plan *plan, int slot, int bins, uint32_t *hist_t, uint32_t *hist_s, double *range4,
                          uint32_t *colour, int64_t *code, double *matched);

/* ---- the reference's individual functions ------------------------------------------------ */
/* rgb_to_lab_pixels / rgb_to_lab_image (LST:67-77, 22-40): uint8 [n,3] -> float64 [n,3] */
int rct_rgb_to_lab(const uint8_t *rgb, double *lab, size_t n, void *stream);
/* lab_to_rgb_pixels / lab_to_rgb_image (LST:80-89, 43-61): float64 [n,3] -> uint8 [n,3];
 * rgb_float (optional, may be NULL) receives the clipped [0,1] doubles before quantisation */
int rct_lab_to_rgb(const double *lab, uint8_t *rgb, double *rgb_float, size_t n, void *stream);
/* apply_pca (LST:92-96 -> sklearn PCA(1).fit_transform, covariance_eigh): float64 [n,3] ->
 * projection [n], mean_[3], components_[3] (device pointers).  |lab| must be < 64.
 * scratch: device memory of rct_pca_scratch_bytes() bytes. */
size_t rct_pca_scratch_bytes(void);
int rct_apply_pca(const double *lab, size_t n, double *proj, double *mean3, double *comp3,
                  void *scratch, void *stream);
/* pca.inverse_transform (LST:161; sklearn _base.py:205): matched [n] -> lab [n,3] */
int rct_pca_inverse(const double *matched, size_t n, const double *mean3, const double *comp3,
                    double *lab, void *stream);
/* match_cdf (LST:99-125): exact sort-based quantile mapping of target_proj [nt] onto
 * source_proj [ns]; matched [nt].  j_index (optional) receives the index np.interp resolves
 * each query to (int64).  scratch: rct_match_scratch_bytes(nt, ns) bytes of device memory. */
size_t rct_match_scratch_bytes(size_t nt, size_t ns);
int rct_match_cdf(const double *target_proj, size_t nt, const double *source_proj, size_t ns,
                  double *matched, int64_t *j_index, void *scratch, void *stream);
/* match_cdf for callers that cannot allocate device memory.  No counterpart in the reference (numpy allocates freely);
 * it exists for hosts where cudaMalloc is off limits at the call site: a plug-in inside an application that owns the
 * whole device heap, a process whose memory pool is sized once at start-up, a real-time loop that cannot afford the
 * device-wide synchronisation an allocation may cause.  Same arguments and same results (bit for bit) as
 * rct_match_cdf, minus `scratch`: the work arrays are a fixed array inside the library image, 34 MiB on every GPU
 * that loads librct.so, held until the process exits.
 *  - limit: nt <= RCT_MATCH_STATIC_MAX and ns <= RCT_MATCH_STATIC_MAX; larger inputs return RCT_ERR_ARG with the
 *    limit in rct_last_error() and touch nothing.
 *  - the call itself performs no cudaMalloc / cudaFree and calls no allocating routine; the caller still provides
 *    target_proj, source_proj, matched and (optionally) j_index in device memory.
 *  - the first call on each device also does one-time set-up (constant tables, kernel attributes, looking up the
 *    array's device address): make that first call during initialisation, before any stream capture begins.
 *  - there is ONE array per device and every call uses it: two calls whose GPU work can overlap (different streams
 *    without an ordering between them, or two host threads) race on it and corrupt each other's result.  Calls on one
 *    stream, or ordered by events, are safe; the array keeps the previous call's contents and the call clears the
 *    parts that must start at zero itself. */
#define RCT_MATCH_STATIC_MAX ((size_t)1 << 20)
int rct_match_cdf_static(const double *target_proj, size_t nt, const double *source_proj, size_t ns,
                         double *matched, int64_t *j_index, void *stream);

/* ---- the steps either side of the hot path --------------------------------
 * f-1: mask-blend tail of AdaIN inference (Style_3DGS/AdaIN/test.py:222-236): mask [mh,mw] float32
 * resized with torch's nearest rule, output [n,3,oh,ow] with torch's bilinear rule
 * (align_corners=False), stylized = content * (1 - mask) + output * mask, planar float32 NCHW.
 * rgb8 (optional) receives what torchvision.utils.save_image quantises, uint8 [n,H,W,3].
 * mask_stride: elements between per-image masks, 0 = one mask for the batch (the reference). */
int rct_adain_mask_blend(const float *content, const float *output, const float *mask, float *stylized,
                         uint8_t *rgb8, int n, int H, int W, int oh, int ow, int mh, int mw,
                         size_t mask_stride, void *stream);
/* f-2: DeepLab post-processing (LST:178-188): softmax over n_classes planes of logits
 * [n_classes, npix] float32, probability of class `cls` > threshold -> uint8 {0,1} */
int rct_softmax_threshold_mask(const float *logits, int n_classes, size_t npix, int cls, float threshold,
                               uint8_t *mask, void *stream);
/* f-2 generalised to label maps (the 8-label configuration): per-pixel arg-max, first index on ties */
int rct_argmax_labels(const float *logits, int n_classes, size_t npix, uint8_t *labels, void *stream);
/* f-4: the temporal-consistency step of the video driver (video/utils.py:355-359).
 * rct_warp_image = warp_image (video/utils.py:89-105): cv2.remap(image, x + flow[0], y + flow[1],
 * INTER_LINEAR, BORDER_REFLECT) in OpenCV's own fixed-point arithmetic; image / warped uint8
 * [n,H,W,channels] (channels 1..4), flow float32 [n,2,H,W].
 * rct_blend_images = blend_images (video/utils.py:223-229) on uint8 inputs: float32
 * alpha*s/255 + (1-alpha)*w/255, *255, clip, truncate; nbytes = number of uint8 elements.
 * rct_warp_blend = blend_images(stylized, warp_image(prev_stylized, flow), alpha) in one pass. */
int rct_warp_image(const uint8_t *image, const float *flow, uint8_t *warped, int n, int H, int W, int channels,
                   void *stream);
int rct_blend_images(const uint8_t *stylized, const uint8_t *warped, double alpha, uint8_t *out, size_t nbytes,
                     void *stream);
int rct_warp_blend(const uint8_t *stylized, const uint8_t *prev_stylized, const float *flow, double alpha,
                   uint8_t *out, int n, int H, int W, int channels, void *stream);

/* f-3: the write at the end of the pipeline, `Image.fromarray(combined_image).save(".../result.jpg")`
 * (LST:241-243; Pillow defaults = libjpeg baseline, 4:2:0, Annex-K Huffman tables, islow DCT), on the device:
 * rgb uint8 [n] frames of H x W x 3, frame_stride bytes apart (0 = packed); out [n][out_stride] receives the
 * complete JFIF files, byte for byte what Pillow writes at the same quality (1..100; the reference uses 75);
 * nbytes[n] (device) = bytes of each file.  A value above out_cap means that file was truncated at out_cap:
 * retry with rct_jpeg_max_bytes(H, W).  workspace: rct_jpeg_workspace_bytes(H, W, n) bytes of device memory,
 * 256-byte aligned, contents irrelevant.  Everything is enqueued on `stream`; nothing synchronises. */
size_t rct_jpeg_workspace_bytes(int H, int W, int n);
size_t rct_jpeg_max_bytes(int H, int W);
int rct_jpeg_encode_async(const uint8_t *rgb, size_t frame_stride, int n, int H, int W, int quality, void *workspace,
                          uint8_t *out, size_t out_stride, size_t out_cap, unsigned long long *nbytes, void *stream);

/* f-3, decode half: `Image.open(path)` of the content / stylised files (LST:200, 219).  Baseline sequential Huffman
 * JPEG, 8-bit, one component or YCbCr at 4:4:4 / 4:2:2 / 4:2:0, with or without restart markers -> the pixels Pillow
 * decodes (libjpeg defaults: islow IDCT, fancy upsampling), bit for bit.  Anything else (progressive, arithmetic,
 * CMYK, 12-bit) returns RCT_ERR_ARG with the reason in rct_last_error(): nothing is decoded approximately.
 * rct_jpeg_decode_info parses the header on the host (no GPU needed).  rct_jpeg_decode_async takes the FILE BYTES IN
 * HOST MEMORY, copies the entropy-coded segment to the device and enqueues the decoder on `stream`; pixels (device)
 * receives uint8 [H,W,3] (or [H,W] for one component).  The host bytes may be released when the call returns.
 * Files with restart markers decode one thread per restart interval. */
/* flags of rct_jpeg_decode_async.  Files WITHOUT restart markers (what Pillow writes by default) are decoded by the
 * self-synchronising parallel Huffman decoder: subsequences of 256 bytes decoded speculatively, their entry states
 * iterated to a fixed point (which proves them exact; the serial kernel takes over on the device if the rounds do not
 * converge) -- same pixels, orders of magnitude faster on large files.  RCT_JPEG_SERIAL forces one thread per restart
 * interval for every file (A/B runs, tests). */
#define RCT_JPEG_SERIAL 1
int rct_jpeg_decode_info(const uint8_t *file_host, size_t nbytes, int *H, int *W, int *channels, size_t *workspace_bytes);
int rct_jpeg_decode_async(const uint8_t *file_host, size_t nbytes, void *workspace, size_t workspace_bytes, uint8_t *pixels,
                          int flags,
                          void *stream);

/* f-3, PNG: `Image.open(path).convert("RGB")` for PNG files (LST:200, 219 take any format Pillow reads).  Decoded ON THE
 * HOST (zlib inflate + scanline filters are serial by construction, see csrc/rct_png.h); the caller uploads the pixels.
 * Non-interlaced, 8-bit grey / grey+alpha / RGB / RGBA and 1..8-bit palette / grey; alpha dropped as convert("RGB") does;
 * anything else returns RCT_ERR_ARG with the reason.  rgb_host receives uint8 [H,W,3]. */
int rct_png_info(const uint8_t *file_host, size_t nbytes, int *H, int *W);
int rct_png_decode_host(const uint8_t *file_host, size_t nbytes, uint8_t *rgb_host, size_t rgb_capacity);

/* f-3 on the end-to-end path: the reference's real boundary is FILES (run_localized_style_transfer(content_img_path, ...)
 * -> ".../localized_style_transfer_result.jpg", LST:191-245).  One call: content file + stylised file (baseline JPEG decoded
 * on the device, PNG decoded on the host; sizes must match the plan's H x W / sh x sw) + the uint8 [H*W] background mask in
 * host memory -> decode -> NEAREST resample -> regional colour transfer -> combine -> baseline JPEG (the bytes Pillow's
 * save(quality) writes) in out_file.  A 4K frame then crosses PCIe as ~3 MB of files + the mask instead of 83 MB of
 * pixels.  Returns the frame status (0 / 1 / 2) or < 0; *out_bytes = size of the file; synchronises.
 * out_capacity: rct_files_out_capacity(H, W) always suffices. */
size_t rct_files_out_capacity(int H, int W);
int rct_transfer_files_host(rct_plan *plan, const uint8_t *content_file, size_t content_bytes, const uint8_t *stylised_file,
                            size_t stylised_bytes, const uint8_t *mask_h, int quality, uint8_t *out_file, size_t out_capacity,
                            size_t *out_bytes, int *status_host);

/* key-table self-check of the last launch group's slot: stats[0] = 1 if the table was built, [1] runs recorded,
 * [2] cells, [3] flagged cells, [4] distinct target colours resolved by a pure cell, [5] of those, colours whose cell
 * disagrees with the colour table (must be 0), [6] colours sent to the table by a flagged cell, [7..9] the same three
 * weighted by pixel count; *max_key_err = max |fp32 key - canonical fp64 key| over all 2^24 colours (bound: 2^-16) */
int rct_plan_debug_keytab(rct_plan *plan, int slot, unsigned long long *stats10, double *max_key_err);

/* ---- introspection for parity tests (RCT_FLAG_DEBUG plans; all synchronise) -------------- */
typedef struct {
    int64_t n_target, n_source;     /* region pixel counts (non-black, LST:134-138) */
    int64_t d_target, d_source;     /* distinct colours per region */
    double t_mean[3], t_comp[3];    /* fg_pca.mean_ / components_[0] */
    double s_mean[3], s_comp[3];    /* bg_pca.* */
    int status;
} rct_frame_info;
int rct_plan_info(rct_plan *plan, int slot, rct_frame_info *info_host);
/* colour-cube histograms: 2 x 2^24 uint32 (target, source), index r | g<<8 | b<<16 */
int rct_plan_debug_hist(rct_plan *plan, int slot, uint32_t *hist_t_host, uint32_t *hist_s_host);
/* per target colour, in ascending key order: colour index, projection key, CDF index j
 * (what np.interp resolved to), matched value, cumulative pixel count.  Arrays of d_target. */
int rct_plan_debug_target(rct_plan *plan, int slot, uint32_t *colour, double *key, int64_t *j,
                          double *matched, uint32_t *cum);
int rct_plan_debug_lut(rct_plan *plan, int slot, uint32_t *lut_host /* 2^24 */);

/* number of kernels launched by this library on this thread since the last reset */
long long rct_launch_count(int reset);

#ifdef __cplusplus
}
#endif
#endif /* RCT_H_ */
