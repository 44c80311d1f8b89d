/*
 * chaqsdf.h -- C ABI of libchaqsdf_b200.so, the B200 (sm_100a) replacement for the hot path of
 * chaquator/chaq-sdfgen's OpenMP program. Plain pointers and sizes only; no C++/torch types.
 * Citations are relative to the reference tree.
 *
 * What each entry point replaces:
 *   dist_transform_2d          openmp/df.h:6 (same symbol, same signature; host pointer, in place)
 *   chaqsdf_u8                 openmp/sdfgen.c:261-300 (transform_img_to_bool :52-62, transform_bool_to_float
 *                              :65-72, dist_transform_2d x2 in OpenMP sections :277-289, transform_float_sub
 *                              :98-106, transform_float_to_byte :75-96) on host buffers
 *   chaqsdf_u8_device          the same, device-resident, stream-ordered, no allocation
 *   chaqsdf_u8_batched_device  the same for a batch of equal-sized images (glyph atlases) in one launch pair
 *   chaqsdf_rowscan_device /   the two passes separately, on row bands, for multi-GPU sharding (the
 *   chaqsdf_envelope_device    reference has no multi-device path)
 *   chaqsdf_fields_device      exact (unwindowed) squared-distance fields + the float fields of
 *                              sdfgen.c:281-294, for parity checks
 *
 * All functions returning int return 0 on success, non-zero on error (chaqsdf_last_error_string()).
 * There is no CPU fallback: without a CUDA device every compute entry fails with CHAQSDF_ERR_CUDA.
 */
#ifndef CHAQSDF_H
#define CHAQSDF_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CHAQSDF_OK 0
#define CHAQSDF_ERR_ARG 1       /* bad argument (zero spread, oversize image, null pointer ...)  */
#define CHAQSDF_ERR_CUDA 2      /* CUDA runtime error or no device                                */
#define CHAQSDF_ERR_WORKSPACE 3 /* caller workspace too small                                     */

#define CHAQSDF_MAX_DIM 32768      /* 15-bit line distances in the 16-bit intermediate            */
#define CHAQSDF_D2_INF 2147483647  /* "no pixel of that class anywhere" in int32 D2 outputs       */

int chaqsdf_version(void);
const char* chaqsdf_last_error_string(void);

/* Number of kernels this library has launched in this process (all threads); for bench accounting. */
unsigned long long chaqsdf_launch_count(void);

/* Tuning switches (run lengths, pipeline band heights, CHAQSDF_DT2D ...; csrc/tuning.h lists them) come from the
 * environment and are read ONCE, when the library is loaded. This re-reads them: a hook for tests and A/B runs. */
void chaqsdf_reload_tuning(void);

/* Per-kernel timing of the banded pass 2 (spread <= 256), for bench.py's roofline: when enabled, the library
 * brackets its own launches with CUDA events on the caller's stream (the first 64 calls after enabling).
 * chaqsdf_profile_mean_ms waits for them and returns the mean milliseconds of { hull build, junction merge,
 * hull paint } and, as its value, the number of calls averaged; on any failure (no profiled call, CUDA error) it
 * returns 0, zeroes ms[] and sets the error string. Not thread-safe; leave it off in production (it adds four
 * event records per call). */
void chaqsdf_profile_enable(int on);
int chaqsdf_profile_mean_ms(float ms[3]);

/* ---- df.h:6 drop-in -------------------------------------------------------------------------------
 * img: host, row-major float[w*h], in place; output sqrtf of the lower envelope min_q |p-q|^2 + img[q], +INF if
 * no sample is finite. Thread-safe (called from two OpenMP section threads at sdfgen.c:277-289). Like the
 * original it has no return value; on error it leaves img untouched and records the error string.
 *   heights all 0 / +INF (what the CLI passes, sdfgen.c:281,286): exact integer kernels; identical to the stock
 *     build wherever that is exact (sides <= 4098), the exact EDT beyond;
 *   any other height, or CHAQSDF_DT2D=stock in the environment at load time: the reference's float32 envelope restated
 *     operation by operation (csrc/dt_float.cuh), bit-identical to the stock build on square images of any size,
 *     including its inexactness beyond side 4098.
 * Works for non-square images (the original does not: df.c:47, df.c:92). */
void dist_transform_2d(float* img, size_t w, size_t h);
int chaqsdf_dt2d_last_status(void); /* status of the calling thread's last dist_transform_2d */

/* ---- fused host-buffer path (sdfgen.c:261-300) ----------------------------------------------------
 * img: host bytes, pixel i at img[i*stride + offset] (stb's (Y,A) buffer: stride 2, offset 1 = alpha,
 * 0 = luminance; sdfgen.c:264). test_above: px > 127 else px < 127 (sdfgen.c:57-59). out: host u8[w*h]. */
int chaqsdf_u8(const uint8_t* img, size_t w, size_t h, size_t stride, size_t offset, int test_above,
               size_t spread, int asymmetric, uint8_t* out);
/* Large images (>= 4 Mpx, spread <= 256) go through chaqsdf_u8 in row bands so that upload, kernels and download
 * overlap. This reports the schedule for a shape: the number of bands (0 = one shot) and, in bands[0..max_bands),
 * their heights in rows, top to bottom. Host-only (no device needed); honours CHAQSDF_PIPE_ROWS / _FIRST. */
size_t chaqsdf_pipeline_plan(size_t w, size_t h, size_t stride, size_t spread, size_t* bands, size_t max_bands);

/* The same over several GPUs of one box, from one process and one host thread: the image is cut
 * into n_devices row bands, band k on devices[k] (a device may be listed more than once). Each device uploads and
 * scans its own rows; the spread rows of the intermediate that pass 2 needs from the neighbouring bands move device
 * to device (cudaMemcpyPeerAsync; NVLink where peer access exists) behind cross-device event waits; the bands' bytes
 * are downloaded into out. Byte-identical to chaqsdf_u8. The caller's current device is restored. */
int chaqsdf_u8_multi(const uint8_t* img, size_t w, size_t h, size_t stride, size_t offset, int test_above,
                     size_t spread, int asymmetric, const int* devices, int n_devices, uint8_t* out);

/* ---- device-resident path -------------------------------------------------------------------------
 * d_img/d_out/d_workspace are device pointers on the current device; stream is a cudaStream_t (NULL =
 * default stream). No allocation or synchronisation inside once the per-(device,spread,asymmetric)
 * quantisation table exists (first call with a new pair builds and uploads it). */
size_t chaqsdf_workspace_bytes(size_t w, size_t h, size_t spread);
int chaqsdf_u8_device(const uint8_t* d_img, size_t w, size_t h, size_t stride, size_t offset, int test_above,
                      size_t spread, int asymmetric, uint8_t* d_out, void* d_workspace, size_t workspace_bytes,
                      void* stream);

/* ---- no-allocation variant ---------------------------------------------------------------------------
 * No counterpart in the reference; added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it). Same computation and
 * bytes as chaqsdf_u8_device, but the workspace AND the quantisation tables live in one file-scope
 * `static __device__` array of CHAQSDF_STATIC_SCRATCH_BYTES inside the library (every GPU that loads the library
 * holds it until exit). The call allocates nothing on the device (no cudaMalloc / cudaFree, no streams or events):
 * it enqueues, on `stream`, an upload of the tables when (spread, asymmetric) differ from the previous call's on
 * this device, then the kernels. Limits: spread <= 256, and
 *   chaqsdf_workspace_bytes(w, h, spread) <= CHAQSDF_STATIC_SCRATCH_BYTES - 180224   (e.g. 2048 x 2048 at spread 256);
 * larger requests fail with CHAQSDF_ERR_WORKSPACE and an error string naming both numbers. All calls on a device
 * share the array, which keeps the previous call's contents (the kernels overwrite what they read; nothing needs
 * zeroing by the caller): calls that overlap on the GPU -- different streams without an ordering between them --
 * race; calls on one stream, or ordered by events, are safe. */
#define CHAQSDF_STATIC_SCRATCH_BYTES 8388608
int chaqsdf_u8_device_static(const uint8_t* d_img, size_t w, size_t h, size_t stride, size_t offset, int test_above,
                             size_t spread, int asymmetric, uint8_t* d_out, void* stream);

/* n images of w*h pixels each, image k at d_imgs + k*w*h*stride, output k at d_out + k*w*h. Images are
 * independent (the batch is treated as separate lines, never as one tall image). */
size_t chaqsdf_batched_workspace_bytes(size_t n, size_t w, size_t h, size_t spread);
int chaqsdf_u8_batched_device(const uint8_t* d_imgs, size_t n, size_t w, size_t h, size_t stride, size_t offset,
                              int test_above, size_t spread, int asymmetric, uint8_t* d_out, void* d_workspace,
                              size_t workspace_bytes, void* stream);

/* ---- the two passes, for row-band sharding --------------------------------------------------------
 * The intermediate between the passes is an opaque buffer of `rows` records of
 * chaqsdf_inter_row_bytes(w, spread) bytes each, one record per image row, rows contiguous (so a run of rows
 * can be handed to another GPU as one block). Its content depends on spread only (not on asymmetric):
 *   spread <= 256: w (rounded up to 128) bytes of min(row distance, 256) - 1, then 1 class bit per pixel;
 *   otherwise    : one uint16 per pixel (bit 15 = class, bits 0..14 = row distance saturated at spread+1).
 * "row distance" = distance along the row to the nearest pixel of the other class.
 * Pass 1 (rows are independent): threshold + row distances for `rows` rows of an image band. */
size_t chaqsdf_inter_row_bytes(size_t w, size_t spread);
int chaqsdf_rowscan_device(const uint8_t* d_img, size_t w, size_t rows, size_t stride, size_t offset,
                           int test_above, size_t spread, void* d_inter, void* stream);
/* Pass 2 (columns; needs `spread` rows of context above and below each output row, or the image edge):
 *   d_inter: buf_rows records, any contiguous run of image rows; output rows [out_begin, out_end) of that
 *   buffer are written to d_out[(r-out_begin)*w + x]. Rows the buffer does not hold are treated as absent,
 *   which is exact when the buffer reaches the image edge or >= spread rows beyond the outputs. */
size_t chaqsdf_envelope_workspace_bytes(size_t w, size_t buf_rows, size_t spread);
int chaqsdf_envelope_device(const void* d_inter, size_t w, size_t buf_rows, size_t out_begin, size_t out_end,
                            size_t spread, int asymmetric, uint8_t* d_out, void* d_workspace,
                            size_t workspace_bytes, void* stream);

/* Pass 2 on the column groups that cover columns [col_begin, col_end) only (spread <= 1024). d_out is still a full-width
 * [out_end-out_begin][w] buffer; only those columns (rounded outwards to multiples of 32, clipped to w) are written.
 * Workspace as chaqsdf_envelope_workspace_bytes(w, buf_rows, spread). For the transposed row-band sharding below. */
int chaqsdf_envelope_cols_device(const void* d_inter, size_t w, size_t buf_rows, size_t out_begin, size_t out_end,
                                 size_t col_begin, size_t col_end, size_t spread, int asymmetric, uint8_t* d_out,
                                 void* d_workspace, size_t workspace_bytes, void* stream);
/* Byte-matrix transposition on the device: dst[c*dst_pitch + r] = src[r*src_pitch + c] for r < rows, c < cols.
 * Transposed row bands ("row bands throughout"): a rank that owns rows [r0, r1) of a big image takes
 * those rows plus `spread` rows of the INPUT on either side, transposes them, and runs pass 1 along the original
 * columns (the only pass that looks across the band's edges, and the cheap one) and pass 2 along the original rows,
 * which are whole inside the band: chaqsdf_rowscan_device on the transposed bytes, chaqsdf_envelope_cols_device on
 * the band's columns, and a transposition back. Replaces nothing in the reference (openmp/df.c:125-136 transposes
 * between its two passes for the same reason: df.c:94). */
int chaqsdf_transpose_u8_device(const uint8_t* d_src, size_t src_pitch, size_t rows, size_t cols, uint8_t* d_dst,
                                size_t dst_pitch, void* stream);

/* ---- glyph / atlas front-end and packer -------------------------------------------------
 * The caller side of a batched glyph job: rasterised glyph canvases in, a packed SDF atlas out. The reference has no
 * counterpart (include/stb/stb_truetype.h and stb_rect_pack.h sit unused in its submodule); csrc/atlas.cu.
 * chaqsdf_atlas_pack (host only, no device needed): skyline bottom-left packing of n rectangles wh[2i] x wh[2i+1] into
 *   an atlas atlas_w pixels wide, `pad` empty pixels right of and below each; writes the top-left corners to
 *   xy[2i], xy[2i+1] and the height used to *atlas_h. Empty rectangles get (0, 0). Deterministic.
 * chaqsdf_glyph_boxes_device: d_boxes[4i..] = {x0, y0, x1, y1} (exclusive), the bounding box of glyph i's inside pixels
 *   (threshold as sdfgen.c:57-59) grown by `margin` and clipped to the canvas; {0,0,0,0} for a canvas without ink.
 * chaqsdf_atlas_blit_device: copies box i of the i-th w*h canvas of d_sdf to d_atlas at (xy[2i], xy[2i+1]).
 * chaqsdf_atlas_u8: all of it from host buffers: upload, boxes (margin = spread), chaqsdf_u8_batched_device, pack,
 *   blit into an atlas cleared to 0 (the byte of a pixel far outside every shape), download of *atlas_h rows;
 *   fails with CHAQSDF_ERR_WORKSPACE when *atlas_h exceeds atlas_rows_cap (then *atlas_h says what is needed). */
int chaqsdf_atlas_pack(size_t n, const uint32_t* wh, size_t atlas_w, size_t pad, uint32_t* xy, size_t* atlas_h);
int chaqsdf_glyph_boxes_device(const uint8_t* d_glyphs, size_t n, size_t w, size_t h, int test_above, size_t margin,
                               int32_t* d_boxes, void* stream);
int chaqsdf_atlas_blit_device(const uint8_t* d_sdf, size_t n, size_t w, size_t h, const int32_t* d_boxes,
                              const uint32_t* d_xy, uint8_t* d_atlas, size_t atlas_w, void* stream);
int chaqsdf_atlas_u8(const uint8_t* glyphs, size_t n, size_t w, size_t h, int test_above, size_t spread, int asymmetric,
                     size_t atlas_w, size_t pad, uint8_t* atlas, size_t atlas_rows_cap, size_t* atlas_h, int32_t* boxes,
                     uint32_t* xy);

/* ---- exact fields, for parity ---------------------------------------------------------------------
 * Unwindowed exact EDT. Any output may be NULL. d2_in/d2_out: int32 squared distance to the nearest
 * inside/outside pixel (CHAQSDF_D2_INF if none). f_in/f_out: sqrtf of those (+INF if none) = the reference's
 * img_float_inside/outside after sdfgen.c:282,287. sdf: the field after sdfgen.c:294. out: the bitmap. */
size_t chaqsdf_fields_workspace_bytes(size_t w, size_t h);
int chaqsdf_fields_device(const uint8_t* d_img, size_t w, size_t h, size_t stride, size_t offset, int test_above,
                          size_t spread, int asymmetric, int32_t* d_d2_in, int32_t* d_d2_out, float* d_f_in,
                          float* d_f_out, float* d_sdf, uint8_t* d_out, void* d_workspace, size_t workspace_bytes,
                          void* stream);

/* ---- exact fields over several GPUs: all-to-all between the passes (SURVEY.md 8(e) variant E1) --------
 * The unwindowed column pass needs whole columns, so the image is cut into row bands for pass 1 and into column
 * parts for pass 2. chaqsdf_col_part gives part q of nparts: columns [*c0, *c1) (parts start at multiples of 32
 * columns; returns the nominal part width). chaqsdf_rowscan_scatter_device runs pass 1 on `rows` rows starting at
 * image row row0 and stores each row straight into the parts' buffers -- d_parts[q] is a device pointer (this
 * device's or a peer-mapped one of another GPU: the all-to-all is the store) to a [h][c1-c0] uint16 buffer holding
 * ALL image rows of that part. After every band has been stored (caller's barrier), each owner runs
 * chaqsdf_fields_from_inter_device on its buffer as an image of width c1-c0: same outputs as chaqsdf_fields_device,
 * for its columns. */
size_t chaqsdf_col_part(size_t w, size_t nparts, size_t q, size_t* c0, size_t* c1);
int chaqsdf_rowscan_scatter_device(const uint8_t* d_img, size_t w, size_t rows, size_t stride, size_t offset,
                                   int test_above, size_t row0, size_t nparts, void* const* d_parts, void* stream);
size_t chaqsdf_fields_from_inter_workspace_bytes(size_t w, size_t h);
int chaqsdf_fields_from_inter_device(const void* d_inter, size_t w, size_t h, size_t spread, int asymmetric,
                                     int32_t* d_d2_in, int32_t* d_d2_out, float* d_f_in, float* d_f_out, float* d_sdf,
                                     uint8_t* d_out, void* d_workspace, size_t workspace_bytes, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* CHAQSDF_H */
