/*
 * rti_b200.h -- C ABI of the B200-native (sm_100a) per-pixel reflectance
 * fitting path of bara96/Smartphone-based-RTI.
 *
 * The reference has no FFI: the path is a set of Python functions in
 * analysis.py that analysis.compute() calls in sequence.  Each entry point
 * below states the reference function / lines whose arithmetic it replaces;
 * the Python host layer (smartphone-based-rti_b200/analysis.py) binds these
 * with ctypes and re-exposes the reference's own call surface.
 *
 * Conventions
 *   - plain pointers and sizes only; no torch / C++ types cross this boundary;
 *   - "d_" pointers are device pointers on the CURRENT CUDA device, "h_"
 *     pointers are host pointers (pinned memory makes the copies asynchronous);
 *   - the caller owns every buffer; the library never frees caller memory;
 *   - device entry points are asynchronous on `stream` (a cudaStream_t passed
 *     as void*; NULL = default stream) and re-entrant across streams/devices;
 *   - return value: 0 = ok, < 0 = bad argument, > 0 = CUDA / numerical error;
 *     rti_last_error_string() describes the last failure on this thread;
 *   - pixel index p = y * cols + x over a row slab [row0, row0 + rows) of the
 *     ROI; y is the GLOBAL ROI row (row0 + local row) because the pixel
 *     position enters the light vector (analysis.py:228-229);
 *   - coefficient maps are planar: coef[k * plane_stride + p], k = 0..5, in the
 *     reference's term order (lu^2, lv^2, lu*lv, lu, lv, 1) (analysis.py:285).
 */
#ifndef RTI_B200_H_
#define RTI_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RTI_OK 0
#define RTI_ERR_ARG (-1)        /* NULL / negative / inconsistent argument                */
#define RTI_ERR_TOO_FEW (-2)    /* PTM with N < 6 (reference: ValueError, analysis.py:296) */
#define RTI_ERR_CUDA 1          /* CUDA runtime error                                      */
#define RTI_ERR_SINGULAR 2      /* >= 1 pixel has a singular system (reference: garbage /
                                   scipy LinAlgError for duplicate RBF nodes)              */
#define RTI_ERR_NONFINITE 3     /* a table value is NaN/inf/out of int32 (reference:
                                   ValueError / OverflowError at analysis.py:407)          */

typedef void *rti_stream_t;     /* cudaStream_t */

/* ---- library ----------------------------------------------------------------- */
int rti_version(void);                       /* 10000*major + 100*minor + patch */
const char *rti_last_error_string(void);
/* SM count / total bytes of HBM of CUDA device `device`. */
int rti_device_info(int device, int *sm_count, int64_t *hbm_bytes);
/* Number of kernels this library has launched on this thread's behalf since the
 * last call with reset != 0 (bench.py's gpu_launches). */
int64_t rti_launch_count(int reset);

/* ---- (1) gather: compute_intensities, analysis.py:196-246 ----------------------
 * d_frames: uint8, points at pixel (x=0, y=row0) of frame 0; the slab of frame i
 * starts at d_frames + i*frame_stride and is row-major [yl][x] (the ROI layout of
 * FeatureMatcher.py:183-184), i.e. slab pixel p = yl*cols + x is byte p.  For a
 * whole-ROI stack [n][R][R], frame_stride = R*R and d_frames = base + row0*cols.
 * d_cams: float64 [n][3] camera positions.
 * Outputs, each [rows*cols][n] with the light index fastest (analysis.py:217-219):
 * d_lx/d_ly float32 = l[0], l[1] of l = (cam - p)/||cam - p|| evaluated in fp64
 * and rounded to float32; d_inten int32. */
int rti_gather_soa(const uint8_t *d_frames, int64_t frame_stride, const double *d_cams,
                   int n, int row0, int rows, int cols,
                   float *d_lx, float *d_ly, int32_t *d_inten, rti_stream_t stream);

/* ---- (2) PTM fit: _interpolate_PTM analysis.py:282-298 -------------------------
 * Least-squares fit of the 6 biquadratic coefficients per pixel.  The reference
 * solves by full SVD; these kernels accumulate the 6x6 normal equations of the
 * same design matrix in fp64 and solve them by an in-register Cholesky, which
 * is the same solution for full-rank pixels.  d_bad (int32, may be NULL) is
 * incremented once per pixel whose system is not positive definite; those
 * pixels' coefficients are NaN.  d_coef: float64 [6][plane_stride]. */

/* fused: straight from the uint8 frames, no SoA materialisation */
int rti_ptm_fit_frames(const uint8_t *d_frames, int64_t frame_stride, const double *d_cams,
                       int n, int row0, int rows, int cols,
                       double *d_coef, int64_t plane_stride, int32_t *d_bad,
                       rti_stream_t stream);
/* per-pixel validity masks (extension; the reference fits every pixel on all frames): a sample
 * takes part in a pixel's fit only when its intensity lies in [lo, hi] (0 <= lo <= hi <= 255), i.e.
 * the result equals _interpolate_PTM (analysis.py:263-298) called with the pixel's masked
 * sub-arrays.  d_valid_count (int32 [n_pixels], may be NULL) receives the number of valid samples;
 * pixels with fewer than 6 (the reference raises there, :296-298) or a singular system are counted
 * in d_bad and get NaN coefficients. */
int rti_ptm_fit_frames_masked(const uint8_t *d_frames, int64_t frame_stride, const double *d_cams,
                              int n, int row0, int rows, int cols, int lo, int hi,
                              double *d_coef, int64_t plane_stride, int32_t *d_valid_count,
                              int32_t *d_bad, rti_stream_t stream);
/* from materialised (lx, ly, intensity) arrays as rti_gather_soa /
 * compute_intensities produce them: one warp per pixel, shuffle reduction */
int rti_ptm_fit_soa(const float *d_lx, const float *d_ly, const int32_t *d_inten,
                    int64_t n_pixels, int n,
                    double *d_coef, int64_t plane_stride, int32_t *d_bad, rti_stream_t stream);
/* shared-light fast path: every pixel sees the same (lx_i, ly_i), so the fit is
 * coef[6][P] = pinv[6][n] . I[n][P] with the 6 x n pseudo-inverse of the common
 * design matrix precomputed by the caller (fp64). */
int rti_ptm_fit_shared(const uint8_t *d_frames, int64_t frame_stride, const double *d_pinv,
                       int n, int64_t n_pixels,
                       double *d_coef, int64_t plane_stride, rti_stream_t stream);

/* Shared-light detection: *d_flag is incremented when some pixel's (lx, ly) row differs
 * (bitwise) from pixel 0's; it stays untouched when all n_pixels rows are identical, i.e.
 * when the data handed to interpolate_intensities (analysis.py:341-343) has one light vector
 * per frame for the whole ROI and the shared-light paths apply. */
int rti_lights_shared(const float *d_lx, const float *d_ly, int64_t n_pixels, int n,
                      int32_t *d_flag, rti_stream_t stream);
/* The 6 x n pseudo-inverse of the common design matrix for rti_ptm_fit_shared, from the
 * shared light components d_lx, d_ly (float32 [n]): pinv = (A^T A)^-1 A^T with A's rows
 * (lu^2, lv^2, lu*lv, lu, lv, 1) formed like analysis.py:284-290; fp64.  d_bad is
 * incremented when A is rank deficient. */
int rti_ptm_shared_pinv(const float *d_lx, const float *d_ly, int n, double *d_pinv,
                        int32_t *d_bad, rti_stream_t stream);

/* ---- (4) light-grid evaluation ----------------------------------------------------
 * d_axis: float64 [g] grid axis values (analysis.py:345-347; the same axis is
 * used for lu and lv).  Value at (v, u): a0*lu^2 + a1*lv^2 + a2*lu*lv + a3*lu +
 * a4*lv + a5 with lu = axis[u], lv = axis[v] (analysis.py:303-313), fp64. */

/* prepare_images_data layout (analysis.py:400-409): d_table int32
 * [g][g][image_stride], image (v, u) holds pixel p at p; the fp64 value is
 * truncated toward zero like the reference's element assignment.  d_bad counts
 * pixels whose values cannot be represented (NaN/inf/|.| >= 2^31). */
int rti_ptm_table(const double *d_coef, int64_t plane_stride, int64_t n_pixels,
                  const double *d_axis, int g,
                  int32_t *d_table, int64_t image_stride, int32_t *d_bad, rti_stream_t stream);
/* interpolate_intensities layout (analysis.py:349, 372): d_grids float64
 * [n_pixels][g][g], pixel-major, indexed [ly][lx]. */
int rti_ptm_grids(const double *d_coef, int64_t plane_stride, int64_t n_pixels,
                  const double *d_axis, int g, double *d_grids, rti_stream_t stream);

/* Fused compute_intensities + PTM fit + table (analysis.py:466-472 for interpolate_PTM=True) in
 * one kernel: every CTA fits 256 pixels and then writes their table entries, so the fp64-bound
 * fit of one CTA overlaps the HBM-bound stores of its neighbours.  Arguments as
 * rti_ptm_fit_frames + rti_ptm_table; d_coef may be NULL when only the table is wanted;
 * d_bad (may be NULL) is int32[2]: [0] singular pixels, [1] pixels with unrepresentable values. */
int rti_ptm_fit_table_frames(const uint8_t *d_frames, int64_t frame_stride, const double *d_cams,
                             int n, int row0, int rows, int cols, const double *d_axis, int g,
                             double *d_coef, int64_t plane_stride,
                             int32_t *d_table, int64_t image_stride, int32_t *d_bad,
                             rti_stream_t stream);

/* prepare_images_data on arbitrary input (analysis.py:400-409): d_grids float64
 * [n_pixels][g][g] -> d_table int32 [g][g][image_stride], truncating toward zero;
 * d_bad counts tiles holding NaN/inf/out-of-int32 values (reference: ValueError /
 * OverflowError at :407).  At most 65535*32 pixels per call. */
int rti_grids_to_table(const double *d_grids, int64_t n_pixels, int g,
                       int32_t *d_table, int64_t image_stride, int32_t *d_bad, rti_stream_t stream);

/* ---- (3) linear RBF: _interpolate_RBF analysis.py:249-260 -> scipy Rbf('linear') --
 * Per pixel: A[j][k] = ||n_j - n_k||_2 over the pixel's n nodes (lx, ly cast to
 * fp64), weights w = solve(A, intensity).  The kernel eliminates node 0 to get
 * the (n-1)x(n-1) symmetric positive definite matrix
 * B[j][k] = |n_j-n_0| + |n_k-n_0| - |n_j-n_k| and Cholesky-factorises it in fp64
 * (no pivoting needed), which yields the same w.  d_work: scratch of
 * rti_rbf_workspace_bytes(n, batch) bytes; pixels are processed `batch` at a
 * time.  d_weights: float64 [n_pixels][n].  d_bad counts singular pixels
 * (duplicate nodes; scipy raises LinAlgError). */
int64_t rti_rbf_workspace_bytes(int n, int batch);
int rti_rbf_fit_soa(const float *d_lx, const float *d_ly, const int32_t *d_inten,
                    int64_t n_pixels, int n, double *d_weights,
                    void *d_work, int64_t work_bytes, int batch,
                    int32_t *d_bad, rti_stream_t stream);
/* rti_rbf_fit_soa without the caller-provided workspace.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA
 * graph capture, tight serving loops, harnesses that forbid allocation).  The Cholesky workspace is
 * a file-scope device array of RTI_RBF_STATIC_SCRATCH_BYTES inside the library, held by every GPU
 * that loads it until the process exits; the call allocates nothing (no cudaMalloc / cudaFree / pool).
 * Same kernel, inputs, outputs and error codes as rti_rbf_fit_soa (d_bad may be NULL); pixels in
 * flight = min(296, scratch / rti_rbf_workspace_bytes(n, 1)).  n whose one-pixel workspace exceeds
 * the scratch (n > 1983) is rejected with RTI_ERR_ARG.  All calls share the scratch: calls that
 * overlap on the GPU (two streams, two host threads on one device) race -- serialise them.  Nothing
 * in the scratch must start at zero; the previous call's contents are overwritten.  The first call on
 * a device looks the array's device address up (cudaGetSymbolAddress): make it outside graph capture. */
#define RTI_RBF_STATIC_SCRATCH_BYTES (32ll << 20)
int rti_rbf_fit_soa_static(const float *d_lx, const float *d_ly, const int32_t *d_inten,
                           int64_t n_pixels, int n, double *d_weights,
                           int32_t *d_bad, rti_stream_t stream);
/* The same fit, tile-preconditioned (rbf_pcg.cu): the SoA arrays describe a rows x cols slab
 * (pixel = y*cols + x).  Per 16x16 pixel tile the reduced matrix of the tile-centre pixel is
 * inverted once in fp64 and serves as the preconditioner of conjugate-gradient solves of every
 * pixel of the tile; matrix entries are regenerated per step in float32 and accumulated in
 * fp64, the preconditioner is applied to all pixels by one batched tcgen05 contraction per
 * step.  Weights reproduce the node values to ~tol (absolute, in intensity units; e.g. 0.01)
 * and scipy's grid values to ~1e-4 of their maximum (north_star allows 1e-3).
 * d_work: rti_rbf_pcg_workspace_bytes(n, tiles) bytes for `tiles` tiles in flight (>= 1; the
 * call batches the slab's tiles through whatever it is given), 256-byte aligned.
 * d_iters: optional int32 [rows*cols], steps each pixel took (negated for a pixel that did
 * not converge).  d_bad: int32[2]; [0] counts
 * tiles whose centre matrix is singular (duplicate nodes), [1] pixels that did not reach tol
 * within max_iters steps OR whose weights are so large (|w|_2 > 10 tol / 1.2e-7: node pairs a few
 * 1e-5 apart) that float32 distances no longer hold the tolerance -- refit those with
 * rti_rbf_fit_soa and evaluate them with rti_rbf_table (fp64).  3 <= n <= 1025. */
int64_t rti_rbf_pcg_workspace_bytes(int n, int64_t tiles);
int rti_rbf_fit_pcg(const float *d_lx, const float *d_ly, const int32_t *d_inten,
                    int rows, int cols, int n, double *d_weights,
                    void *d_work, int64_t work_bytes, int max_iters, double tol,
                    int32_t *d_iters, int32_t *d_bad, rti_stream_t stream);
/* Rbf.__call__ on the grid: value(v,u) = sum_k ||(axis[u], axis[v]) - n_k|| w_k.
 * Table layout / truncation / d_bad as rti_ptm_table. */
int rti_rbf_table(const float *d_lx, const float *d_ly, const double *d_weights,
                  int64_t n_pixels, int n, const double *d_axis, int g,
                  int32_t *d_table, int64_t image_stride, int32_t *d_bad, rti_stream_t stream);
/* The same table with the node distances formed in float32 (one MUFU.SQRT per term) and the
 * products with the fp64 weights and the running sums kept in fp64.  Absolute error of a grid
 * value ~ 1.2e-7 * |w|_2 (each distance carries a float32 rounding): ~5e-5 of the largest value
 * for typical pixels (north_star allows 1e-3), but over budget for pixels with node pairs ~1e-5
 * apart, whose weights reach 1e6 -- rti_rbf_fit_pcg reports those pixels, evaluate them with
 * rti_rbf_table.  An int32 entry can differ from the reference's, by one, only where the fp64
 * value lies that close to an integer.  Bound by the SFU (G*G*n square roots per pixel),
 * ~1.8x the fp64 form. */
int rti_rbf_table_fast(const float *d_lx, const float *d_ly, const double *d_weights,
                       int64_t n_pixels, int n, const double *d_axis, int g,
                       int32_t *d_table, int64_t image_stride, int32_t *d_bad, rti_stream_t stream);
/* fp64 grids, layout as rti_ptm_grids */
int rti_rbf_grids(const float *d_lx, const float *d_ly, const double *d_weights,
                  int64_t n_pixels, int n, const double *d_axis, int g,
                  double *d_grids, rti_stream_t stream);

/* ---- (3b) linear RBF, shared-light case ----------------------------------------------
 * When every pixel sees the same n nodes (lx_i, ly_i) -- pixels_lx[y][x][:] identical for
 * all (y, x) in the data handed to interpolate_intensities (analysis.py:341-343, 361) -- the
 * per-pixel `Rbf(...)` + `rbf(xi, yi)` of analysis.py:259-260 is one matrix A for all
 * pixels, so
 *     grid values [g*g][P] = (Phi_grid . A^-1)[g*g][n] . D[n][P]      (RTI_RBF_OP_GRID)
 *     weights     [n][P]   =  A^-1[n][n]               . D[n][P]      (RTI_RBF_OP_WEIGHTS)
 * with D the intensities.  The operator is built once in fp64 by the per-pixel kernels run on
 * n unit-intensity pseudo-pixels, each row scaled by a power of two and split into fp16
 * hi + lo; the contraction runs on the tcgen05 tensor cores (TMA-fed, fp32 accumulation in
 * TMEM; intensities are exact in fp16).
 *
 * d_op_hi / d_op_lo: fp16 bit patterns, rti_rbf_shared_operator_elems(op_rows, n) elements
 * each ([op_rows rounded up to 128][n rounded up to 64], lights fastest), 16-byte aligned;
 * d_inv_scale: float [op_rows] (op_rows = g*g or n).  d_bad counts singular systems
 * (duplicate nodes; scipy: LinAlgError).
 * d_packed: rti_rbf_shared_packed_elems(n, n_pixels) fp16 elements, 16-byte aligned: pixel
 * blocks of 256, each a contiguous [n][256] slab (element (light i, pixel p) at
 * ((p / 256) * n + i) * 256 + p % 256), produced by one of the two pack calls: from the uint8 frames
 * (layout of rti_gather_soa's d_frames) or from int32 intensities [n_pixels][n]
 * (compute_intensities layout; values outside [-2048, 2048] are not exact in fp16 and are
 * counted in d_bad). */
#define RTI_RBF_OP_GRID 0
#define RTI_RBF_OP_WEIGHTS 1
int64_t rti_rbf_shared_operator_elems(int op_rows, int n);
int64_t rti_rbf_shared_packed_elems(int n, int64_t n_pixels);
int rti_rbf_shared_operator(const float *d_lx, const float *d_ly, int n, const double *d_axis, int g,
                            int kind, uint16_t *d_op_hi, uint16_t *d_op_lo, float *d_inv_scale,
                            int32_t *d_bad, rti_stream_t stream);
int rti_rbf_shared_pack_frames(const uint8_t *d_frames, int64_t frame_stride, int n, int64_t n_pixels,
                               uint16_t *d_packed, rti_stream_t stream);
int rti_rbf_shared_pack_soa(const int32_t *d_inten, int64_t n_pixels, int n, uint16_t *d_packed,
                            int32_t *d_bad, rti_stream_t stream);
/* prepare_images_data layout: d_table int32 [op_rows][image_stride] (op_rows = g*g), values
 * truncated toward zero; d_bad counts threads that saw a value outside int32. */
int rti_rbf_shared_table(const uint16_t *d_packed, int n, int64_t n_pixels, const uint16_t *d_op_hi,
                         const uint16_t *d_op_lo, const float *d_inv_scale, int op_rows,
                         int32_t *d_table, int64_t image_stride, int32_t *d_bad, rti_stream_t stream);
/* pixel-major float64 [n_pixels][op_rows]: the layout of rti_rbf_grids (grid operator) or of
 * rti_rbf_fit_soa's d_weights (weight operator). */
int rti_rbf_shared_apply_f64(const uint16_t *d_packed, int n, int64_t n_pixels, const uint16_t *d_op_hi,
                             const uint16_t *d_op_lo, const float *d_inv_scale, int op_rows,
                             double *d_out, rti_stream_t stream);

/* ---- relighting: relighting_event interactive_relighting.py:25-38 -----------------
 * Re-render one light direction (lu, lv) from the coefficient map: value =
 * trunc(poly(lu, lv)) clamped to [0, 255] (analysis.py:404-407 +
 * interactive_relighting.py:35-36).  d_hsv: uint8 [n_pixels][3] HSV ROI or NULL.
 * With d_hsv the value replaces V and the pixel is converted with OpenCV's 8-bit
 * COLOR_HSV2BGR arithmetic into d_out uint8 [n_pixels][3]; with d_hsv == NULL
 * d_out is uint8 [n_pixels] (the V plane only).
 * row_width: columns of the ROI image (n_pixels = rows * row_width).  cv2.cvtColor converts
 * row by row and its scalar tail -- the last row_width % 32 pixels of every row on an AVX2
 * host -- rounds x*255 half to even where the vector body truncates; the kernels reproduce
 * that per pixel column.  row_width == 0: vector-body arithmetic for every pixel. */
int rti_relight_ptm(const double *d_coef, int64_t plane_stride, int64_t n_pixels, int row_width,
                    double lu, double lv, const uint8_t *d_hsv, uint8_t *d_out,
                    rti_stream_t stream);
/* same from a float32 coefficient map (24 B/pixel; the fp32 evaluation can differ
 * from the table by 1 where the fp64 value is within ~1e-4 of an integer) */
int rti_relight_ptm_f32(const float *d_coef, int64_t plane_stride, int64_t n_pixels, int row_width,
                        float lu, float lv, const uint8_t *d_hsv, uint8_t *d_out,
                        rti_stream_t stream);
/* float64 -> float32 planar coefficient copy for rti_relight_ptm_f32 */
int rti_coef_to_f32(const double *d_coef, int64_t n_values, float *d_out, rti_stream_t stream);
/* table lookup variant: clamp an int32 table image to [0,255] and convert */
int rti_relight_from_image(const int32_t *d_image, int64_t n_pixels, int row_width, const uint8_t *d_hsv,
                           uint8_t *d_out, rti_stream_t stream);

/* ---- ROI extraction upstream of the path ---------------------------
 * get_ROI(static_img, points, hsv=True) (Utils/utilities.py:302-342): crop
 * [y_min, y_min+rows) x [x_min, x_min+cols) out of n_frames BGR frames
 * uint8 [height][width][3] (frame_stride bytes apart) and convert with OpenCV's
 * 8-bit COLOR_BGR2HSV integer arithmetic.  hsv != 0: d_out uint8 [rows][cols][3]
 * per frame (the relighter's HSV ROI, interactive_relighting.py:109); hsv == 0:
 * only V = max(B,G,R), uint8 [rows][cols] per frame = one frame of the stack that
 * compute_intensities reads (FeatureMatcher.py:183-184).  out_stride in bytes.
 * The ROI bounds (centre, clamping quirks) are host arithmetic, see roi.py.
 * Any byte alignment of d_bgr, the frame rows and d_out is accepted.  Rows are read in whole
 * 16-byte granules: up to 15 bytes before a run (never before the 16-byte granule that holds
 * d_bgr[0]) and after it (never past the last frame's end) are touched but not used. */
int rti_roi_extract(const uint8_t *d_bgr, int64_t frame_stride, int n_frames, int height,
                    int width, int x_min, int y_min, int rows, int cols, int hsv,
                    uint8_t *d_out, int64_t out_stride, rti_stream_t stream);

/* ---- host-buffer pipeline: analysis.compute() steps 2-4 for PTM ---------------------
 * analysis.py:466-472 in one call on HOST buffers: frames uint8 [n][rows_total*cols]
 * and cams float64 [n][3] go in, the slab [row0, row0+rows) is fitted and evaluated,
 * and h_table int32 [g][g][rows*cols] (prepare_images_data layout) and/or h_coef
 * float64 [6][rows*cols] come back (either may be NULL).  Row tiles are streamed:
 * H2D of tile t+1, fit + table of tile t and D2H of tile t-1 overlap on three
 * streams.  Uses the current device; `tile_rows` <= 0 picks a default.  Blocks
 * until the result is in host memory.  The cached workspace (streams, tile buffers) is
 * per device ordinal: calls for different devices run concurrently (one host thread per
 * GPU), calls for the same device serialise on that device's workspace.
 * Narrow transfer: PCIe is the bound of this call (4 B per pixel and light direction).
 * When the process has the host to itself (no LOCAL_WORLD_SIZE > 1) and >= 8 cores, a
 * tile crosses PCIe as int16 (narrowed on the device, which also checks that every value
 * fits; a tile with a value outside int16 is copied again as int32) into a pinned staging
 * ring, and host threads sign-extend it into h_table while the next tile is in flight:
 * same table, half the PCIe bytes, and h_table need not be pinned.
 * RTI_HOST_TRANSFER=wide|narrow forces either form (mixed: int16, but a tile goes as int32
 * straight into a pinned h_table when the staging ring is full; measured slower on a
 * 16-core host), RTI_HOST_THREADS the thread count (default: the affinity mask /
 * LOCAL_WORLD_SIZE, at most 16). */
int rti_ptm_compute_host(const uint8_t *h_frames, const double *h_cams, int n,
                         int rows_total, int cols, int row0, int rows,
                         const double *h_axis, int g,
                         int32_t *h_table, double *h_coef, int tile_rows);
/* One tile of a DEVICE table into a HOST table through the same staging ring and host threads
 * (what prepare_images_data, analysis.py:375-411, needs for any interpolation mode):
 * d_tile int32 [n_images][image_stride], of which tile_pixels per image are used, goes to
 * h_table[i * host_image_stride + pix_off + x] (pinned or pageable).  The copy is ordered behind
 * `ready_event` (a cudaEvent_t; the producer's completion) or, when that is NULL, behind
 * everything queued on `stream` so far; it blocks until the data is in h_table.  int16 or
 * int32 crossing as for rti_ptm_compute_host (int32 when a value of the tile is outside int16). */
int rti_table_tile_to_host(const int32_t *d_tile, int64_t image_stride, int64_t n_images,
                           int64_t tile_pixels, int32_t *h_table, int64_t host_image_stride,
                           int64_t pix_off, void *ready_event, rti_stream_t stream);
/* The same into a FILE: image i's values go to byte h_offsets[i] + 4 * pix_off of the open file
 * descriptor fd, written with pwrite by the host threads -- the image payloads of the result
 * pickle that ut.write_on_file(.., compressed=False) produces (Utils/utilities.py:48-70,
 * analysis.py:475), laid out by table_io.TableTileWriter.  h_offsets: host int64 [n_images]. */
int rti_table_tile_to_file(const int32_t *d_tile, int64_t image_stride, int64_t n_images,
                           int64_t tile_pixels, int fd, const int64_t *h_offsets, int64_t pix_off,
                           void *ready_event, rti_stream_t stream);
/* bytes the calling thread's last rti_ptm_compute_host / rti_table_tile_to_host moved device -> host (bench.py's
 * e2e.d2h_bytes_per_step: half of the table's size when it crossed as int16) */
int64_t rti_host_d2h_bytes(void);
/* ---- frame-list ingest helpers (host only, no CUDA) -----------------------------------
 * ut.read_from_file (Utils/utilities.py:84-96) decodes '<name>.pbz2' with one thread; ingest.py
 * decodes the file's bzip2 blocks in parallel and uses these two byte loops for what would
 * otherwise stay serial Python: finding the 48-bit block magic at any BIT offset (ascending bit
 * offsets into bit_offsets[0, cap); the return value is the number found, which may exceed cap;
 * threads <= 0: one per core) and copying a bit range (MSB first within a byte; returns the
 * bit offset after the copy; the tail of the last byte is zeroed). */
int64_t rti_find_magic48(const uint8_t *buf, int64_t n, uint64_t magic48, int64_t *bit_offsets,
                         int64_t cap, int threads);
int64_t rti_copy_bits(const uint8_t *src, int64_t src_bit0, int64_t nbits, uint8_t *dst,
                      int64_t dst_bit0);
/* release the cached device workspace / streams of rti_ptm_compute_host */
int rti_destroy(void);

#ifdef __cplusplus
}
#endif
#endif /* RTI_B200_H_ */
