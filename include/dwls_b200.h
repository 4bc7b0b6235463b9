/* dwls_b200.h — C-ABI of the B200-native disparity WLS filtering library (libdwls_b200.so).
 *
 * This is the drop-in boundary: the entry points below are what a binding of the reference's
 * ximgproc disparity-filtering path would call instead of the CPU implementation.  Plain
 * pointers, sizes and strides only (no OpenCV, no torch types).  All file:line citations are
 * relative to the reference tree (modules/ximgproc/...).
 *
 *   DisparityWLSFilterImpl::filter        src/disparity_filters.cpp:219-298   -> dwls_filter_host / dwls_filter_batch_dev
 *   DisparityWLSFilter getters            src/disparity_filters.cpp:126-139   -> conf / roi_out arguments
 *   FastGlobalSmootherFilterImpl::init    src/fgs_filter.cpp:141-173          -> dwls_fgs_* (weights are rebuilt per call,
 *   FastGlobalSmootherFilterImpl::filter  src/fgs_filter.cpp:182-233             LUT cached per sigma inside the ctx)
 *   fastGlobalSmootherFilter              src/fgs_filter.cpp:687-691          -> dwls_fgs_host / dwls_fgs_dev
 *   ComputeDiscontinuityAwareLRC_ParBody  src/disparity_filters.cpp:306-341   -> dwls_lrc_mask_host (debug / parity export)
 *
 * Error model: every call returns a dwls_status; nothing throws across the ABI.  The
 * reference raises cv::Exception through CV_Assert / CV_Error (disparity_filters.cpp:221-222,
 * 262-264; fgs_filter.cpp:143-144, 184-189); the host facades map DWLS_BAD_ARG / DWLS_BAD_SIZE
 * back to exceptions.  There is NO CPU fallback: without a CUDA device every compute entry
 * point fails with DWLS_CUDA_ERROR.
 *
 * Threading: one ctx = one CUDA device + its streams + one workspace; calls on one ctx from different CUDA streams are ordered
 * by the library (each call waits for the previous call's work), they do not run concurrently; use one ctx per host
 * thread (the reference's filter objects are not re-entrant either, fgs_filter.cpp:198-223).
 */
#ifndef DWLS_B200_H
#define DWLS_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(_WIN32)
#define DWLS_API __declspec(dllexport)
#else
#define DWLS_API __attribute__((visibility("default")))
#endif

typedef enum dwls_status {
    DWLS_OK = 0,
    DWLS_BAD_ARG = 1,     /* CV_Assert on depth / channels / emptiness / parameter range */
    DWLS_BAD_SIZE = 2,    /* CV_Error(StsBadSize) src vs guide size; ROI outside the image   */
    DWLS_CUDA_ERROR = 3,  /* no device, launch or allocation failure                         */
    DWLS_UNSUPPORTED = 4  /* inside the API but outside what the kernels cover (documented)  */
} dwls_status;

/* depth codes are OpenCV's: CV_8U = 0, CV_16S = 3, CV_32F = 5 */
enum { DWLS_8U = 0, DWLS_16S = 3, DWLS_32F = 5 };

/* image descriptor = the fields of a cv::Mat header the path reads */
typedef struct dwls_image {
    void*   data;      /* first pixel (host or device depending on the entry point) */
    int64_t step;      /* bytes between rows (cv::Mat::step)                         */
    int     rows, cols;
    int     depth;     /* DWLS_8U / DWLS_16S / DWLS_32F                              */
    int     channels;
} dwls_image;

/* DisparityWLSFilterImpl state (disparity_filters.cpp:52-63, defaults 142-159, 215) */
typedef struct dwls_params {
    double lambda;                /* 8000.0 */
    double sigma_color;           /* 1.0    */
    int    lrc_thresh;            /* 24     */
    int    disc_radius;           /* 5      */
    int    left_offset, right_offset, top_offset, bottom_offset;   /* matcher-derived margins, 401/407 */
    int    use_confidence;        /* createDisparityWLSFilterGeneric(bool) */
    double lambda_attenuation;    /* 0.25 (edge_filter.hpp:393 default)    */
    int    num_iter;              /* 3                                      */
} dwls_params;

typedef struct dwls_ctx dwls_ctx;

DWLS_API void        dwls_default_params(dwls_params* p);
DWLS_API dwls_status dwls_ctx_create(int device, dwls_ctx** out);
DWLS_API void        dwls_ctx_destroy(dwls_ctx* ctx);
DWLS_API const char* dwls_last_error(const dwls_ctx* ctx);   /* ctx may be NULL: last create error */
DWLS_API const char* dwls_version(void);
/* number of kernels the ctx has launched so far (bench.py reports the per-step delta) */
DWLS_API int64_t     dwls_kernel_launches(const dwls_ctx* ctx);
/* per-kernel CUDA-event timing on the launching stream (off by default).  The report synchronises
 * the device, returns one entry per kernel that ran since the last report and resets the record. */
DWLS_API dwls_status dwls_set_kernel_timing(dwls_ctx* ctx, int on);
DWLS_API int         dwls_kernel_timing_report(dwls_ctx* ctx, const char** names, int64_t* launches, double* total_ms, int cap);
/* CUDA-graph replay of a filter call's launch sequence (on by default).  A sequence is captured the second time
 * the same buffers, geometry and parameters are passed and replayed as one graph launch from then on; kernels counted
 * by dwls_kernel_launches include the replayed ones.  Timing mode always uses plain launches. */
DWLS_API dwls_status dwls_set_cuda_graphs(dwls_ctx* ctx, int on);
DWLS_API int64_t     dwls_graph_replays(const dwls_ctx* ctx);
/* bytes the host-buffer entry points have moved over PCIe since the context was created (either pointer may be NULL).
 * dwls_filter_batch_host uploads only the part of each input the filter reads: the ROI of the left disparity map and of
 * the view, the mirrored ROI of the right disparity map (disparity_filters.cpp:202-203), each with a few pixels of margin. */
DWLS_API void        dwls_transfer_bytes(const dwls_ctx* ctx, int64_t* h2d, int64_t* d2h);
/* layout of the two solver planes of the confidence path between the sweeps: off (the default) = two row-major planes;
 * on = one array interleaved sector by sector, so that a column strip of both planes is 64 contiguous bytes per row
 * (faster column pass, slower row pass: measured a net loss).  Results are identical either way; the
 * switch exists for A/B measurements and tests. */
DWLS_API dwls_status dwls_set_plane_interleave(dwls_ctx* ctx, int on);
/* dwls_filter_batch_dev deals its launch groups round-robin to `lanes` (1..4) internal streams forked from / joined into
 * the caller's stream, each lane with its own workspace slice (default 1: every group on the caller's stream). */
DWLS_API dwls_status dwls_set_lanes(dwls_ctx* ctx, int lanes);
/* frames processed per kernel launch group in the batch entry point (0 = choose automatically) */
DWLS_API dwls_status dwls_set_group_frames(dwls_ctx* ctx, int frames);

/* ---- DisparityWLSFilter::filter on HOST buffers (what the cv::Mat facade calls) -------------
 * disp_left/right: CV_16SC1, same size; view: CV_8UC1/3; roi_in: user ROI {x,y,w,h} in
 * disparity-map coordinates or NULL / zero area to derive it from the offsets (228-233).
 * out16s: CV_16SC1 at view size, caller-allocated.  conf32f: optional CV_32FC1 at view size
 * (getConfidenceMap, 138/274).  roi_out: the disparity-resolution ROI used (getROI, 139).
 * Copies H2D, runs the kernels, copies D2H, synchronises.  The buffers may be ordinary pageable memory: transfers of more
 * than 2 MB are staged through pinned memory by four worker threads the context keeps (started on first use, parked between calls) (each with its own copy
 * stream); smaller ones take the driver's pageable path.  dwls_fgs_host moves its images the same way. */
DWLS_API dwls_status dwls_filter_host(dwls_ctx* ctx,
                                      const dwls_image* disp_left, const dwls_image* disp_right,
                                      const dwls_image* view, const int* roi_in, const dwls_params* params,
                                      dwls_image* out16s, dwls_image* conf32f, int* roi_out);

/* ---- same, on DEVICE buffers, for a batch of independent frames ------------------------------
 * Frame f of every array starts at base + f * frame_stride (bytes).  All frames share sizes,
 * ROI and parameters.  Asynchronous on `stream` (a cudaStream_t; NULL = the ctx stream). */
typedef struct dwls_batch {
    int n_frames;
    const int16_t* disp_left;  int64_t dl_step, dl_frame;    /* bytes */
    const int16_t* disp_right; int64_t dr_step, dr_frame;    /* NULL when !use_confidence */
    int disp_cols, disp_rows;
    const uint8_t* view;       int64_t view_step, view_frame;
    int view_cols, view_rows, view_channels;
    int16_t* out;              int64_t out_step, out_frame;   /* view size */
    float*   conf;             int64_t conf_step, conf_frame; /* view size; NULL to skip the export */
} dwls_batch;

DWLS_API dwls_status dwls_filter_batch_dev(dwls_ctx* ctx, const dwls_batch* batch, const int* roi_in,
                                           const dwls_params* params, int* roi_out, void* stream);

/* Same batch, but every pointer of `batch` is HOST memory (pinned memory gives full PCIe speed).
 * The library pipelines H2D copies, kernels and D2H copies over two streams and returns when the
 * host outputs are complete. */
DWLS_API dwls_status dwls_filter_batch_host(dwls_ctx* ctx, const dwls_batch* batch, const int* roi_in,
                                            const dwls_params* params, int* roi_out);

/* ---- fastGlobalSmootherFilter ---------------------------------------------------------------
 * guide CV_8UC1/3; src/dst CV_8U/16S/32F with 1..4 channels, same size as guide (else
 * DWLS_BAD_SIZE, fgs_filter.cpp:185-189). */
DWLS_API dwls_status dwls_fgs_host(dwls_ctx* ctx, const dwls_image* guide, const dwls_image* src, dwls_image* dst,
                                   double lambda, double sigma_color, double lambda_attenuation, int num_iter);
DWLS_API dwls_status dwls_fgs_dev(dwls_ctx* ctx, const dwls_image* guide, const dwls_image* src, dwls_image* dst,
                                  double lambda, double sigma_color, double lambda_attenuation, int num_iter,
                                  void* stream);

/* ---- fastGlobalSmootherFilter on device buffers WITHOUT any allocation -------------------------------------------
 * No counterpart in the reference: added for callers that cannot call cudaMalloc
 * (code under CUDA-graph capture, tight serving loops, harnesses that forbid it).  Same arguments, kernels and results as
 * dwls_fgs_dev, but the weight LUT, the two weight planes and the solver planes live in ONE static device array of
 * DWLS_NOALLOC_SCRATCH_BYTES (16 MiB) inside the library - every GPU that loads the library holds it until the process
 * exits - and the call makes no cudaMalloc / cudaFree / event or stream creation.
 * Limit: 786432 + (2 + channels) * round_up(cols, 8) * rows * 4 bytes must fit (dwls_fgs_noalloc_required_bytes; e.g. CV_32FC1
 * up to 1280 x 720, 8UC3 up to 960 x 540), at most 2176 rows; larger images return DWLS_UNSUPPORTED with the sizes in
 * dwls_last_error.  Per-kernel timing must be off.
 * All calls (of all contexts on a device) share the array, which keeps the last call's contents: each call clears the plane
 * region it uses and rebuilds the LUT, so calls issued one after another on a stream - or ordered by events - are
 * independent; calls that OVERLAP on the GPU race on the array and give undefined results. */
#define DWLS_NOALLOC_SCRATCH_BYTES (16u << 20)
DWLS_API dwls_status dwls_fgs_dev_noalloc(dwls_ctx* ctx, const dwls_image* guide, const dwls_image* src, dwls_image* dst,
                                          double lambda, double sigma_color, double lambda_attenuation, int num_iter,
                                          void* stream);
DWLS_API size_t      dwls_fgs_noalloc_scratch_bytes(void);
DWLS_API size_t      dwls_fgs_noalloc_required_bytes(int cols, int rows, int channels);

/* ---- peer-memory gather of a sharded batch (one process per GPU, one NVLink / NVSwitch box) --------------------------
 * No counterpart in the reference (its path is single-process); this is the exchange step of SURVEY section 8e.  The gather
 * root allocates the buffer of the whole batch (dwls_peer_alloc: cudaMalloc + 64-byte CUDA IPC handle, which the caller
 * transports to the other ranks), they map it (dwls_peer_open) and push every finished chunk of results to its final place
 * with a peer-to-peer copy on the context's copy stream (dwls_peer_push: ordered after the work queued on `after_stream`,
 * returns at once; copy engines over NVLink, no SM involved).  dwls_peer_flush makes a stream wait for all pushes issued so
 * far; the caller then synchronises the ranks (any stream-ordered collective) before the root reads the buffer. */
DWLS_API dwls_status dwls_peer_alloc(dwls_ctx* ctx, size_t bytes, void** dev_ptr, unsigned char* handle64);
DWLS_API dwls_status dwls_peer_free(dwls_ctx* ctx, void* dev_ptr);
DWLS_API dwls_status dwls_peer_open(dwls_ctx* ctx, const unsigned char* handle64, void** mapped);
DWLS_API dwls_status dwls_peer_close(dwls_ctx* ctx, void* mapped);
DWLS_API dwls_status dwls_peer_push(dwls_ctx* ctx, void* dst_mapped, const void* src, size_t bytes, void* after_stream);
DWLS_API dwls_status dwls_peer_flush(dwls_ctx* ctx, void* stream);
DWLS_API int64_t     dwls_peer_bytes(const dwls_ctx* ctx);

/* ---- parity / debug exports ------------------------------------------------------------------
 * LRC decision per disparity pixel: 0 outside the ROI, 1 right index outside the right ROI
 * (confidence kept), 2 consistent (min of both confidences), 3 inconsistent (confidence 0). */
DWLS_API dwls_status dwls_lrc_mask_host(dwls_ctx* ctx, const dwls_image* disp_left, const dwls_image* disp_right,
                                        int view_cols, const int* roi_in, const dwls_params* params,
                                        uint8_t* mask, int64_t mask_step);
/* weight planes Chor/Cvert for a host guide (fgs_filter.cpp:593-661), dense w*h floats each */
DWLS_API dwls_status dwls_fgs_weights_host(dwls_ctx* ctx, const dwls_image* guide, double sigma_color,
                                           float* chor, float* cvert);

/* ---- stereo block matching on the device: the producer of the filter's inputs ---------------------
 * Replaces cv::StereoBM::compute as the reference's sample calls it (samples/disparity_filtering.cpp:148-160) for the
 * matcher pair createDisparityWLSFilter / createRightMatcher configure (src/disparity_filters.cpp:389-403, 419-428):
 * PREFILTER_XSOBEL with cap, SAD over a block_size^2 window, winner-takes-all, optional texture / uniqueness tests,
 * parabola sub-pixel fit, FILTERED = (min_disparity - 1) * 16 outside the valid-disparity rectangle.  disp12MaxDiff and
 * speckle filtering are not implemented (both factories switch them off, :389-390, :426-427).  Bit-exact CV_16S. */
typedef struct dwls_bm_params {
    int num_disparities;      /* positive, divisible by 16 */
    int block_size;           /* odd, >= 5 */
    int min_disparity;
    int pre_filter_cap;       /* 31 */
    int texture_threshold;    /* cv default 10; createDisparityWLSFilter sets 0 */
    int uniqueness_ratio;     /* cv default 15; createDisparityWLSFilter sets 0 */
} dwls_bm_params;

/* a batch of stereo views on the device (frame f of an array at base + f * frame stride, bytes) */
typedef struct dwls_views {
    int n_frames;
    const uint8_t* left;  int64_t left_step, left_frame;
    const uint8_t* right; int64_t right_step, right_frame;
    int cols, rows, channels;      /* CV_8UC1, or CV_8UC3 in BGR order (converted like cvtColor BGR2GRAY) */
} dwls_views;

DWLS_API void        dwls_bm_default_params(dwls_bm_params* p);
/* disp_left = matcher.compute(left, right); disp_right (NULL to skip) = createRightMatcher(matcher).compute(right, left),
 * i.e. the same matcher on swapped views with minDisparity -(min + num) + 1.  Strides in bytes.  Asynchronous on `stream`. */
DWLS_API dwls_status dwls_stereo_bm_dev(dwls_ctx* ctx, const dwls_views* views, const dwls_bm_params* params,
                                        int16_t* disp_left, int64_t dl_step, int64_t dl_frame,
                                        int16_t* disp_right, int64_t dr_step, int64_t dr_frame, void* stream);
/* same for one pair of host images; disp_right may be NULL */
DWLS_API dwls_status dwls_stereo_bm_host(dwls_ctx* ctx, const dwls_image* left, const dwls_image* right, const dwls_bm_params* params,
                                         dwls_image* disp_left, dwls_image* disp_right);
/* the sample's downscale step on the device (cv::resize by 0.5 of even-sized views, samples/disparity_filtering.cpp:138-139) */
DWLS_API dwls_status dwls_halve_views_dev(dwls_ctx* ctx, const dwls_views* views, uint8_t* left_out, uint8_t* right_out,
                                          int64_t out_step, int64_t out_frame, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* DWLS_B200_H */
