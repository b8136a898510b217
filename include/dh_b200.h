/*
 * dh_b200.h — C ABI of the B200-native (sm_100a) Gaussian heatmap renderer.
 *
 * This is the drop-in boundary for ONE hot path of ACCV-Lab: accvlab.draw_heatmap.
 * Every entry point below replaces a C++-linkage launcher (or a torch-op chain) of the
 * reference; citations are relative to /root/reference/packages/draw_heatmap ("DH").
 *
 * Conventions
 *   - plain pointers and sizes only; no torch / ATen / C++ types cross this boundary;
 *   - all data pointers are DEVICE pointers unless the name ends in `_host`;
 *   - `stream` is a `cudaStream_t` passed as `void*` (NULL = legacy default stream);
 *   - every call is asynchronous on `stream`, allocates no memory and keeps no state (exceptions: cached
 *     device attributes, and the two helper streams of the host-buffer entry);
 *   - return value: DH_OK (0), a negative DH_ERR_* code, or a positive `cudaError_t`
 *     reported by the launch (the reference only printf()s launch errors,
 *     DH/accvlab/draw_heatmap/csrc/draw_heatmap_cuda.cu:37-40,56-59; we return them);
 *   - there is NO CPU fallback: without a CUDA device every compute entry point fails.
 *
 * Arithmetic contract (fp32, IEEE round-to-nearest, no fast-math), restating
 * DH/accvlab/draw_heatmap/include/draw_heatmap_cuda_kernel.cuh:36-49,61-68,90-97:
 *     d = 2r+1; sigma = (float)d / factor; var = 2.0f*sigma*sigma; vinv = 1.0f/var
 *     window: rows [y-r, y+r] x cols [x-r, x+r], clipped to the map (square, not a disc)
 *     v(i,j) = expf(-((float)(i*i) + (float)(j*j)) * vinv) * k_scale
 *     heat[p][y+i][x+j] = (heat >= v) ? heat : v
 */
#ifndef DH_B200_H
#define DH_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DH_B200_VERSION 100 /* 0.1.0 */

/* status codes (negative = ours, positive = cudaError_t) */
#define DH_OK 0
#define DH_ERR_INVALID_ARGUMENT (-1)
#define DH_ERR_WORKSPACE_TOO_SMALL (-2)
#define DH_ERR_UNSUPPORTED_SHAPE (-3)
#define DH_ERR_NO_DEVICE (-4)
#define DH_ERR_STATIC_SCRATCH_EXCEEDED (-5) /* dh_draw_heatmap_batched_noalloc: problem beyond DH_NOALLOC_MAX_SCRATCH_BYTES */

/* blend mode */
#define DH_MODE_ACCUMULATE 0 /* reference semantics: max-blend into existing contents (read+write touched pixels) */
#define DH_MODE_FRESH 1      /* zero-init fused: every pixel of every plane is written exactly once, nothing is read */

/* dtype tag of the per-sample target counts (RaggedBatch.sample_sizes) */
#define DH_SIZES_I32 0
#define DH_SIZES_I64 1

/* Limits checked by every entry point (DH_ERR_UNSUPPORTED_SHAPE otherwise):
 *   1 <= height, width;  height*width < 2^31;  width <= 2^23;
 *   objects with radius < 0 draw nothing (as in the reference: empty loops);
 *   objects with radius > DH_MAX_RADIUS are skipped and counted in dh_status.skipped_radius. */
#define DH_MAX_RADIUS 4194304

/* Optional device-side status record (16 bytes, caller-zeroed, may be NULL).
 * The reference traps on a bad label with a device assert (cuh:102) and never checks
 * heatmap_idxes (cuh:59,70 — out-of-bounds write). We never write out of bounds: such
 * objects are skipped and counted here. */
typedef struct dh_status {
    uint32_t skipped_label;  /* label outside [0, num_classes) */
    uint32_t skipped_plane;  /* heatmap_idx outside [0, num_heatmaps) */
    uint32_t skipped_radius; /* radius > DH_MAX_RADIUS */
    uint32_t reserved;
} dh_status;

/* ---- introspection ------------------------------------------------------------------ */
int dh_version(void);
/* static string: compiler, arch (sm_100a), build flags */
const char* dh_build_info(void);
/* static string for any value returned by the functions below */
const char* dh_status_string(int status);
/* number of kernels launched by calls of this library in the calling process (for bench gpu_launches) */
uint64_t dh_launch_count(void);

/* ---- K1: bbox -> integer centre + radius ---------------------------------------------
 * Replaces the torch-op chain get_centers_and_radii (DH/tests/_test_helpers.py:20-28):
 *     radius = max(1, (int32) ceilf(min(cx-x1, cy-y1, x2-cx, y2-cy) / osf))
 *     xi = (int32)(cx / osf); yi = (int32)(cy / osf)        (IEEE division, truncation)
 * centers_xy  f32 [n,2] (x,y);  bboxes_xyxy f32 [n,4] (x1,y1,x2,y2)
 * centers_out i32 [n,2];        radii_out   i32 [n]
 * Padded RaggedBatch slots are computed like any other slot (the reference does the same). */
int dh_radius_center(void* stream, const float* centers_xy, const float* bboxes_xyxy, float out_size_factor,
                     int64_t n, int32_t* centers_out, int32_t* radii_out);

/* ---- batched entry (RaggedBatch layout) ----------------------------------------------
 * Alignment: `centers` (and `centers_xy` below) should be 8-byte and `bboxes_xyxy` 16-byte aligned — each slot is then
 * read with one vector load; arrays that are only 4-byte aligned (a contiguous torch view at an odd storage offset)
 * are accepted and read with scalar loads, as the reference does (cuh:57-58,86-87).
 * Replaces draw_heatmap_batched_cuda(cudaStream_t, float*, const int*, const int*, const int*,
 *   int H, int W, int T, float factor, float k, int B, int C = 0, const int* labels = nullptr)
 * (DH/accvlab/draw_heatmap/csrc/draw_heatmap_cuda.cu:43-60; kernel cuh:76-109).
 *   heatmap      f32 [B,H,W] (labels == NULL) or [B,C,H,W] (labels != NULL), contiguous
 *   centers      i32 [B,T,2] (x,y);  radii i32 [B,T];  labels i32 [B,T] or NULL
 *   nums_targets [B], int32 or int64 per `nums_dtype`; values are clamped to [0,T]
 *   num_classes  C; ignored (treated as 1) when labels == NULL
 *   mode         DH_MODE_ACCUMULATE | DH_MODE_FRESH
 *   status       optional device dh_status*, may be NULL
 * Plane offsets are 64-bit (the reference overflows int32 at B*C*H*W >= 2^31, cuh:99-104).
 * Note: `labels == NULL` selects the one-plane-per-sample layout. A class-wise call with max_num_targets == 0 (empty
 * labels array) must still pass a non-NULL pointer (it is never dereferenced). */
int dh_draw_heatmap_batched(void* stream, float* heatmap, const int32_t* centers, const int32_t* radii,
                            const void* nums_targets, int nums_dtype, const int32_t* labels, int64_t batch_size,
                            int64_t max_num_targets, int64_t num_classes, int64_t height, int64_t width,
                            float diameter_to_sigma_factor, float k_scale, int mode, dh_status* status);

/* K1 fused in: same as dh_radius_center followed by dh_draw_heatmap_batched, one launch.
 *   centers_xy f32 [B,T,2], bboxes_xyxy f32 [B,T,4]; optional outputs centers_out i32 [B,T,2] and
 *   radii_out i32 [B,T] (both NULL or both non-NULL; all B*T slots are written, padded ones included). */
int dh_draw_heatmap_batched_from_boxes(void* stream, float* heatmap, const float* centers_xy,
                                       const float* bboxes_xyxy, float out_size_factor, const void* nums_targets,
                                       int nums_dtype, const int32_t* labels, int64_t batch_size,
                                       int64_t max_num_targets, int64_t num_classes, int64_t height,
                                       int64_t width, float diameter_to_sigma_factor, float k_scale, int mode,
                                       int32_t* centers_out, int32_t* radii_out, dh_status* status);

/* dh_draw_heatmap_batched_from_boxes in DH_MODE_FRESH with the result digest FUSED into the render launch: while
 * every touched pixel is finalised the kernel accumulates dh_checksum's three sums (see below; untouched pixels are
 * +0.0f and contribute nothing) and adds them to `digest_out3` (DEVICE, 3 x uint64, zeroed or carried over by the
 * caller) — the heatmap is not read back from HBM for it. `index_base` is the element index of heatmap[0] in the
 * unsharded result (rank offset when a batch is split over GPUs). Equals dh_checksum(heatmap) bit for bit.
 * Returns DH_ERR_UNSUPPORTED_SHAPE for shapes the main kernel does not serve (width % 4 != 0, width > 1024 or a
 * heatmap base that is not 16-byte aligned): call the plain entry + dh_checksum then. */
int dh_draw_heatmap_batched_from_boxes_digest(void* stream, float* heatmap, const float* centers_xy,
                                              const float* bboxes_xyxy, float out_size_factor,
                                              const void* nums_targets, int nums_dtype, const int32_t* labels,
                                              int64_t batch_size, int64_t max_num_targets, int64_t num_classes,
                                              int64_t height, int64_t width, float diameter_to_sigma_factor,
                                              float k_scale, int64_t index_base, uint64_t* digest_out3,
                                              dh_status* status);

/* ---- the same entries with a caller-provided bin-list workspace (K2 as a pre-pass) -----------------------
 * SURVEY.md 8(b) "C-level ABI": "a caller-provided workspace pointer + size (plus a *_workspace_bytes query) for the
 * bin lists". With a DEVICE workspace of at least dh_batched_workspace_bytes(B, T, C) bytes (16-byte aligned; contents
 * undefined before and after) the call is TWO launches: bin_kernel reads every slot of the RaggedBatch arrays once
 * (validity cuh:84, label, K1 for float boxes, window / sigma arithmetic cuh:61-68) and counting-sorts each sample's
 * objects by class into compact per-plane lists; the renderer then takes a plane's list with one 8-byte load per unit
 * and renders planes of <= 32 objects warp-autonomously (no block barrier). Results are bit-identical to the entries
 * above; dh_status counts every skipped object exactly once. workspace == NULL (or too small, or a shape the main
 * renderer does not serve, or num_classes > 2048) falls back to the single-launch path of the entry above — never an
 * error. The workspace must not be shared by calls that may run concurrently. */
size_t dh_batched_workspace_bytes(int64_t batch_size, int64_t max_num_targets, int64_t num_classes);
int dh_draw_heatmap_batched_ws(void* stream, float* heatmap, const int32_t* centers, const int32_t* radii,
                               const void* nums_targets, int nums_dtype, const int32_t* labels, int64_t batch_size,
                               int64_t max_num_targets, int64_t num_classes, int64_t height, int64_t width,
                               float diameter_to_sigma_factor, float k_scale, int mode, dh_status* status,
                               void* workspace, size_t workspace_bytes);
int dh_draw_heatmap_batched_from_boxes_ws(void* stream, float* heatmap, const float* centers_xy,
                                          const float* bboxes_xyxy, float out_size_factor, const void* nums_targets,
                                          int nums_dtype, const int32_t* labels, int64_t batch_size,
                                          int64_t max_num_targets, int64_t num_classes, int64_t height,
                                          int64_t width, float diameter_to_sigma_factor, float k_scale, int mode,
                                          int32_t* centers_out, int32_t* radii_out, dh_status* status,
                                          void* workspace, size_t workspace_bytes);
int dh_draw_heatmap_batched_from_boxes_digest_ws(void* stream, float* heatmap, const float* centers_xy,
                                                 const float* bboxes_xyxy, float out_size_factor,
                                                 const void* nums_targets, int nums_dtype, const int32_t* labels,
                                                 int64_t batch_size, int64_t max_num_targets, int64_t num_classes,
                                                 int64_t height, int64_t width, float diameter_to_sigma_factor,
                                                 float k_scale, int64_t index_base, uint64_t* digest_out3,
                                                 dh_status* status, void* workspace, size_t workspace_bytes);

/* ---- allocation-free variant of dh_draw_heatmap_batched_ws ------------------------------------------------
 * NO COUNTERPART IN THE REFERENCE: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid allocation). Same arguments,
 * same results (bit for bit) as dh_draw_heatmap_batched / dh_draw_heatmap_batched_ws; the bin-list workspace of the
 * two-launch path is a file-scope static __device__ array inside the library (DH_NOALLOC_MAX_SCRATCH_BYTES = 8 MiB;
 * every GPU that loads the library holds that much device memory until the process exits). The call allocates nothing
 * (no cudaMalloc / cudaFree / allocating library routine); the caller passes inputs and outputs only.
 *   limit:   dh_batched_workspace_bytes(B, T, C) <= DH_NOALLOC_MAX_SCRATCH_BYTES (e.g. B*T <= ~520 k slots at C = 10;
 *            cfg5, B = 3072 x T = 100 x C = 10, needs 5.2 MB), else DH_ERR_STATIC_SCRATCH_EXCEEDED — never a silent
 *            change of path. Shapes the two-launch path does not serve (see the *_ws entries) run the single-launch
 *            renderer, which needs no scratch at all.
 *   sharing: ONE array per device, shared by every call in the process, keeping the last call's contents; nothing in
 *            it has to start at zero (every record and reachable list entry is rewritten by the call's first launch
 *            before its second reads it). Calls that may OVERLAP on the GPU (different streams or threads) race on it:
 *            serialise them, or use the *_ws entry with one workspace each. */
#define DH_NOALLOC_MAX_SCRATCH_BYTES (8u << 20)
size_t dh_noalloc_max_scratch_bytes(void);
int dh_draw_heatmap_batched_noalloc(void* stream, float* heatmap, const int32_t* centers, const int32_t* radii,
                                    const void* nums_targets, int nums_dtype, const int32_t* labels,
                                    int64_t batch_size, int64_t max_num_targets, int64_t num_classes, int64_t height,
                                    int64_t width, float diameter_to_sigma_factor, float k_scale, int mode,
                                    dh_status* status);

/* ---- 16-bit heatmap planes (not in the reference, whose heatmaps are float32 only:
 * draw_heatmap_cuda.cu:84,117,160 `data_ptr<float>()`) -------------------------------------
 * DH_MODE_FRESH render into fp16 or bf16 planes: every value is computed in fp32 exactly as by the fp32 entries and
 * rounded once (to nearest even) when it is stored, so the result equals the fp32 result cast to the 16-bit type,
 * bit for bit; half the output bytes. Objects come EITHER from (centers i32 [B,T,2], radii i32 [B,T]) OR from float
 * boxes (centers_xy f32 [B,T,2], bboxes_xyxy f32 [B,T,4], out_size_factor) — pass the other pair as NULL.
 * Shapes: width % 4 == 0, width <= 1024, (height*width) % 8 == 0, heatmap base 16-byte aligned; anything else
 * returns DH_ERR_UNSUPPORTED_SHAPE (render fp32 and convert). No accumulate mode, no fused digest. */
#define DH_OUT_F16 1
#define DH_OUT_BF16 2
int dh_render_heatmap_batched_16(void* stream, void* heatmap, int out_dtype, const int32_t* centers,
                                 const int32_t* radii, const float* centers_xy, const float* bboxes_xyxy,
                                 float out_size_factor, const void* nums_targets, int nums_dtype,
                                 const int32_t* labels, int64_t batch_size, int64_t max_num_targets,
                                 int64_t num_classes, int64_t height, int64_t width, float diameter_to_sigma_factor,
                                 float k_scale, dh_status* status);
/* ... with the bin-list workspace of dh_batched_workspace_bytes (see above) */
int dh_render_heatmap_batched_16_ws(void* stream, void* heatmap, int out_dtype, const int32_t* centers,
                                    const int32_t* radii, const float* centers_xy, const float* bboxes_xyxy,
                                    float out_size_factor, const void* nums_targets, int nums_dtype,
                                    const int32_t* labels, int64_t batch_size, int64_t max_num_targets,
                                    int64_t num_classes, int64_t height, int64_t width, float diameter_to_sigma_factor,
                                    float k_scale, dh_status* status, void* workspace, size_t workspace_bytes);

/* ---- flat entry (concatenated objects + per-object plane index) ----------------------
 * Replaces draw_heatmap_cuda(cudaStream_t, float*, const int*, const int*, const int*, int H, int W,
 *   float factor, float k, int N, int M) (draw_heatmap_cuda.cu:29-41; kernel cuh:51-74).
 *   heatmaps f32 [M,H,W]; centers i32 [N,2]; radii i32 [N]; heatmap_idxes i32 [N]
 *   workspace: device scratch of at least dh_draw_heatmap_workspace_bytes(N, M) bytes, 16-byte aligned,
 *              used for the object->plane bin lists (counting sort); contents undefined afterwards. */
size_t dh_draw_heatmap_workspace_bytes(int64_t num_targets, int64_t num_heatmaps);
int dh_draw_heatmap(void* stream, float* heatmaps, const int32_t* centers, const int32_t* radii,
                    const int32_t* heatmap_idxes, int64_t num_targets, int64_t num_heatmaps, int64_t height,
                    int64_t width, float diameter_to_sigma_factor, float k_scale, int mode, void* workspace,
                    size_t workspace_bytes, dh_status* status);

/* ---- result checksum (multi-GPU verification and the e2e "result read") -----------------
 * Order-independent, exactly additive digest of `n` fp32 values (so shards of a batch can be digested on
 * different GPUs and summed with wrapping uint64 adds, e.g. an NCCL all_reduce(SUM) over int64):
 *     out[0] += sum_i  bits(v_i)                          (fp32 bit patterns, wrapping uint64)
 *     out[1] += sum_{i: v_i > 0} mix(index_base + i)      (hash of the touched-pixel mask; mix = 32-bit
 *                                                          murmur3-finaliser of the index, widened to 64 bits)
 *     out[2] += #{i: v_i > 0}
 * `out` is a DEVICE array of 3 uint64 that the caller zeroes (or carries over). No reference equivalent:
 * the reference has no multi-GPU path. */
int dh_checksum(void* stream, const float* values, int64_t n, int64_t index_base, uint64_t* out3);

/* ---- host-buffer entry (end-to-end measurement and non-torch callers) -----------------
 * Inputs live in HOST memory (pinned for full speed). The call stages the inputs to the device with
 * cudaMemcpyAsync, renders in FRESH mode with K1 fused (dh_draw_heatmap_batched_from_boxes), and
 *   - heatmap_host != NULL: copies the heatmap back to the host in chunks of `chunk_samples` samples; render of
 *     chunk i overlaps the device->host copy of chunk i-1 on two internal (per-device, lazily created,
 *     process-lifetime) streams that are fork/joined to `stream` with events. If heatmap_device is NULL the
 *     chunks are rendered into two staging buffers inside `device_scratch`, so device memory use is bounded
 *     independent of B; otherwise they are rendered in place into heatmap_device [B,(C),H,W].
 *   - heatmap_host == NULL: heatmap_device must be non-NULL and the result stays on the device. When the inputs
 *     are large (>= 1.5 MiB) they are cut into 2-6 sample chunks that grow geometrically (x4) from a first one of
 *     ~100 KiB of inputs — only the first copy is exposed — or into up to 16 equal chunks of `chunk_samples` samples
 *     when > 0: chunk c+1 crosses PCIe on an internal copy stream while chunk c is rendered, consecutive chunks
 *     alternating between the two internal render streams; otherwise (or with chunk_samples < 0) the inputs are
 *     brought over and rendered by ONE launch on `stream` — when all of them lie in pinned (device-mapped) host memory
 *     and are <= 8 MiB by one gather launch that reads them in place over PCIe, else by one cudaMemcpyAsync per array.
 *   - checksum_host (optional, 3 x uint64, host): dh_checksum of the whole result, copied to the host (computed
 *     inside the render launch when the shape allows it, by a separate dh_checksum pass otherwise).
 * labels_host == NULL -> heatmap is [B,H,W] and num_classes is ignored. nums_targets_host is int64 [B].
 * chunk_samples == 0 selects a default (host output: about 64 MiB of output per chunk); < 0 disables input pipelining.
 * `device_scratch` (16-byte aligned) must hold dh_host_scratch_bytes(...) bytes for the same arguments.
 * The call is asynchronous with respect to the host like every other entry: synchronise `stream` before
 * reading heatmap_host / checksum_host. */
size_t dh_host_scratch_bytes(int64_t batch_size, int64_t max_num_targets, int64_t num_classes, int64_t height,
                             int64_t width, int has_labels, int64_t chunk_samples, int staging_buffers);
int dh_draw_heatmap_batched_from_boxes_host(void* stream, float* heatmap_host, float* heatmap_device,
                                            uint64_t* checksum_host, const float* centers_xy_host,
                                            const float* bboxes_xyxy_host, float out_size_factor,
                                            const int64_t* nums_targets_host, const int32_t* labels_host,
                                            int64_t batch_size, int64_t max_num_targets, int64_t num_classes,
                                            int64_t height, int64_t width, float diameter_to_sigma_factor,
                                            float k_scale, void* device_scratch, size_t device_scratch_bytes,
                                            int64_t chunk_samples);

/* ---- N1: list of per-sample arrays -> RaggedBatch layout (the step before the renderer) ----------------
 * Replaces the per-sample slice-assign loop of combine_data
 * (/root/reference/packages/batching_helpers/accvlab/batching_helpers/batched_processing_py.py:411-423) and of
 * /root/reference/packages/draw_heatmap/examples/draw_heatmap_batched.py:42-50 by ONE launch.
 * dst [B, T, row_elems] elements of `elem_bytes` (1,2,4,8,16) bytes; sample b takes its first
 * n_b = clamp(sizes[b], 0, T) rows from src_ptrs[b] (row-major contiguous, aligned to elem_bytes) and ZEROS after
 * (the reference starts from torch.zeros). `src_ptrs` (B device pointers) and `sizes` (int64 [B]) are DEVICE
 * arrays; src_ptrs[b] is not dereferenced when n_b == 0. Every destination element is written exactly once. */
int dh_pack_ragged(void* stream, void* dst, const void* const* src_ptrs, const int64_t* sizes, int64_t batch_size,
                   int64_t max_sample_size, int64_t row_elems, int elem_bytes);

/* Same with K1 fused (dh_radius_center arithmetic): per-sample float centres [n_b,2] (8-byte aligned) and boxes
 * [n_b,4] (16-byte aligned) -> centers_out int32 [B,T,2] (8-byte aligned), radii_out int32 [B,T]; padded slots = 0. */
int dh_pack_radius_center(void* stream, const float* const* centers_ptrs, const float* const* bboxes_ptrs,
                          const int64_t* sizes, float out_size_factor, int64_t batch_size, int64_t max_sample_size,
                          int32_t* centers_out, int32_t* radii_out);

/* ---- N2: DALI `draw_gaussians` operator semantics (second renderer mode) --------------------------------
 * GPU counterpart of the reference's CPU-only DALI operator
 * (/root/reference/packages/dali_pipeline_framework/ext_impl/DrawGaussians.cc:31-75 draw_gaussian, :88-157 RunImpl):
 *   out = in; for every slot t < nums_targets[b] with active[b,t]: plane = slice_ids[b,t], k = k_for_classes[plane],
 *   window cols [floor((float)cx - r), ceil((float)cx + r)] x rows likewise (skipped when wholly outside, clipped),
 *   sigma = r * radius_to_sigma_factor, val = k * expf(-(float)(dy^2 + dx^2) / (2 sigma^2)), pixel = max(pixel, val).
 * Layout: RaggedBatch instead of DALI's per-sample tensor lists — centers int32 [B,T,2] (x,y; 8-byte aligned),
 * radii float32 [B,T], slice_ids int32 [B,T] (NULL -> 0), active uint8/bool [B,T] (NULL -> all active),
 * nums_targets [B] (tag nums_dtype), k_for_classes float32 [num_k] DEVICE array, heatmaps float32 [B,C,H,W]
 * ([B,H,W]: pass C = 1). heatmap_in == NULL: FRESH (out = max(0, ...), every pixel written once, nothing read);
 * heatmap_in == heatmap_out: in place (touched pixels only); otherwise out-of-place like the operator.
 * Deliberate differences: a slice_id outside [0, min(C, num_k)) skips the object and counts it in
 * dh_status.skipped_label (the reference fails the batch, or writes out of bounds when slice_id >= C); radii that are
 * NaN or |r| > 2048 are skipped and counted in skipped_radius (the reference's int32 squares overflow / its
 * float -> int casts are undefined there). One launch (three with the workspace below). */
int dh_draw_gaussians(void* stream, const float* heatmap_in, float* heatmap_out, const int32_t* centers,
                      const float* radii, const int32_t* slice_ids, const uint8_t* active, const void* nums_targets,
                      int nums_dtype, const float* k_for_classes, int64_t num_k, int64_t batch_size,
                      int64_t max_num_targets, int64_t num_classes, int64_t height, int64_t width,
                      float radius_to_sigma_factor, dh_status* status, void* workspace, size_t workspace_bytes);
/* Optional DEVICE workspace (16-byte aligned, dh_draw_gaussians_workspace_bytes(B, T) bytes) enables the FRESH fast
 * path (heatmap_in == NULL, output >= 256 MiB, width % 4 == 0, width <= 1024, status == NULL): 3 launches — classify the objects; the
 * main renderer (dh_draw_heatmap_batched's kernel, bulk stores + cluster launch control) draws every object whose
 * window is the symmetric integer square; the exact kernel then redraws, in place, only the objects it is not
 * (fp32-rounded asymmetric windows, negative radii, degenerate sigma) and returns at once when there are none.
 * workspace == NULL: the single exact kernel. Results are identical either way. */
size_t dh_draw_gaussians_workspace_bytes(int64_t batch_size, int64_t max_num_targets);

/* Front half of the reference's BoundingBoxToHeatmapConverter — everything between the annotation boxes and
 * draw_gaussians, ~10 DALI CPU operators per sample in the reference — as ONE launch over the RaggedBatch slots:
 *   centre from the box (numba_operators.py:811-815, when centers == NULL), scale by scale_xy[b] = (sx, sy) and clip to
 *   the heatmap (python_operator_functions.py:140-171), integer centre = floor, sub-pixel offset (converter :352-355),
 *   is_active = class < num_classes && clipped (h,w) >= min sizes && area fraction >= min_fraction_area && is_valid
 *   (python_operator_functions.py:213-251, converter :369-371), radius = clamp(max(0, min distance of the clipped centre
 *   to the clipped box border) * radius_scaling_factor [float64 product stored to fp32], min_radius, max_radius)
 *   (numba_operators.py:865-904, converter :376).
 * Inputs (device): bboxes f32 [B,T,4] (x1,y1,x2,y2 in image coordinates, corners in any order; 16-byte aligned),
 * centers f32 [B,T,2] or NULL, classes i32 [B,T] or NULL (no class check), is_valid u8 [B,T] or NULL,
 * nums_targets [B], scale_xy f32 [B,2] (heatmap size / image size per sample), min_sizes_hw f64 [num_min_sizes,2]
 * (h,w): per-class table indexed by class when classes != NULL, else its row 0; num_min_sizes = 0: no size check.
 * Outputs (device, [B,T,...], padded slots = 0): center_int i32 [.,2], radii f32, is_active u8 (inputs of
 * dh_draw_gaussians), and optionally bboxes_clipped f32 [.,4], center_offset f32 [.,2], hw_clipped f32 [.,2] (h,w),
 * fraction_area f32 (NULL to skip). NaN coordinates are not supported (numpy's clip propagates them). */
int dh_boxes_to_gaussian_params(void* stream, const float* bboxes, const float* centers, const int32_t* classes,
                                const uint8_t* is_valid, const void* nums_targets, int nums_dtype, const float* scale_xy,
                                const double* min_sizes_hw, int64_t num_min_sizes, int64_t num_classes,
                                int64_t batch_size, int64_t max_num_targets, int64_t heatmap_height,
                                int64_t heatmap_width, double radius_scaling_factor, float min_radius, float max_radius,
                                double min_fraction_area, float* bboxes_clipped, int32_t* center_int,
                                float* center_offset, float* radii, uint8_t* is_active, float* hw_clipped,
                                float* fraction_area);

/* ---- N4: CenterNet / CenterPoint gaussian_radius(det_size, min_overlap) as a K1 mode ----------------------
 * NOT in the reference (its helper uses the minimum border distance, dh_radius_center): PARITY UNPINNED. Restates the
 * published algorithm in float64 like upstream: DH_RADIUS_CORNERNET = CornerNet / CenterNet utils/image.py / CenterPoint
 * center_utils.py (roots (b + sqrt(b^2 - 4ac)) / 2), DH_RADIUS_MMDET = mmdet gaussian_target.py (corrected roots
 * (b -/+ sqrt) / (2a)); radii_out[i] = max(min_radius, (int)min(r1, r2, r3)). det_hw float32 [n,2] = (height, width)
 * in heatmap pixels; radii_float_out (optional) receives min(r1, r2, r3) before truncation. */
#define DH_RADIUS_CORNERNET 0
#define DH_RADIUS_MMDET 1
int dh_gaussian_radius(void* stream, const float* det_hw, double min_overlap, int variant, int32_t min_radius, int64_t n,
                       int32_t* radii_out, float* radii_float_out);

/* ---- N3: fused Gaussian-focal-loss consumer (the step AFTER the renderer) ------------------------------
 * loss(pred, target) of /root/reference/packages/draw_heatmap/docs/intro.rst:7-26 where `target` is the heatmap
 * dh_draw_heatmap_batched(FRESH) would draw from the same RaggedBatch inputs — but the target is never written to
 * HBM: each pixel's value is formed in registers and consumed at once.
 *   target == 1 : loss = -(1 - pred)^alpha * log(pred + eps)
 *   otherwise   : loss = -pred^alpha * log(1 - pred + eps) * (1 - target)^gamma          (eps as in mmdet, may be 0)
 * pred float32 [B,(C),H,W] in (0,1); grad_pred (optional, same shape) receives d loss / d pred per pixel (unscaled);
 * loss_out2 = DEVICE double[2]: { sum of the loss over all pixels, number of pixels with target == 1 } — summed in a
 * fixed order (per-unit partials in `workspace`, then one block), so results are run-to-run identical.
 * Accuracy: logarithm <= 1.05 ulp (own polynomial, see csrc/dh_gaussians.cu log_pos; arguments below 2^-126 count as
 * 0), gradient quotient 2 ulp (__fdividef). fast_math != 0 (alpha = 2, gamma = 4 only): logarithm through the
 * special-function unit (__logf; ~1e-6 relative on the sum). grad_pred == NULL selects forward-only kernels.
 * k_scale > 0; radii in [0, 2048] (others draw nothing). The reference documents the formula but has no implementation:
 * PARITY UNPINNED (oracle = float64 numpy restatement, tests/test_focal_loss.py). Two launches. */
size_t dh_focal_loss_workspace_bytes(int64_t batch_size, int64_t num_classes, int64_t height, int64_t width);
int dh_gaussian_focal_loss(void* stream, const float* pred, const int32_t* centers, const int32_t* radii,
                           const void* nums_targets, int nums_dtype, const int32_t* labels, int64_t batch_size,
                           int64_t max_num_targets, int64_t num_classes, int64_t height, int64_t width,
                           float diameter_to_sigma_factor, float k_scale, float alpha, float gamma, float eps,
                           int fast_math, float* grad_pred, double* loss_out2, void* workspace,
                           size_t workspace_bytes);

/* Backward of dh_gaussian_focal_loss: dst[i] = src[i] * (*scale_device) for n float32 elements, the scalar read from
 * DEVICE memory (the upstream gradient of the scalar loss as autograd hands it over; no host synchronisation). dst may
 * equal src. One streaming launch at copy rate. */
int dh_scale_by_device_scalar(void* stream, float* dst, const float* src, const float* scale_device, int64_t n);

/* ---- measurement helper ---------------------------------------------------------------
 * Pure 128-bit-store fill of `bytes` bytes (multiple of 16, 16-byte aligned) with zeros using the same
 * store path and grid shape as the FRESH renderer: the write-only ceiling the renderer is compared to. */
int dh_fill_probe(void* stream, void* dst, size_t bytes);

/* out[i] = expf(x_i) where x_i is the float whose bit pattern is start_bits + i, evaluated with exactly the
 * exponential the renderer uses. Lets the tests check on the GPU (exhaustively over the argument range) that it
 * is monotone — the property behind "one expf per pixel": max_o k*exp(a_o) == k*exp(max_o a_o) — and measure its
 * ulp distance to a correctly rounded exp. */
int dh_expf_probe(void* stream, uint32_t start_bits, int64_t n, float* out);

#ifdef __cplusplus
}
#endif
#endif /* DH_B200_H */
