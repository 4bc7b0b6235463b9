// dwls_capi.cu — C-ABI of libdwls_b200.so (declared in include/dwls_b200.h): context, workspace,
// kernel orchestration for DisparityWLSFilter::filter and fastGlobalSmootherFilter.
// Reference orchestration being replaced: modules/ximgproc/src/disparity_filters.cpp:219-298 and
// modules/ximgproc/src/fgs_filter.cpp:141-233.  There is no CPU path in this file.
#include <cuda_runtime.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <functional>
#include <algorithm>
#include <atomic>
#include <mutex>
#include <condition_variable>
#include <string>
#include <thread>
#include <vector>
#include <unordered_map>
#include <unordered_set>

#include "../../include/dwls_b200.h"
#include "dwls_resize.cuh"
#include "dwls_stereo.cuh"

using namespace dwls;

constexpr int DWLS_MAX_LANES = 4;

// persistent host workers of a context (staging copies of the single-call host entry points): started on first use,
// parked on a condition variable between calls, joined when the context is destroyed
class WorkerPool {
    std::vector<std::thread> th_;
    std::mutex m_;
    std::condition_variable cv_, done_;
    const std::function<void(int)>* job_ = nullptr;
    uint64_t gen_ = 0;
    int pending_ = 0;
    bool stop_ = false;
    void loop(int id)
    {
        uint64_t seen = 0;
        for (;;) {
            const std::function<void(int)>* job;
            {
                std::unique_lock<std::mutex> lk(m_);
                cv_.wait(lk, [&] { return stop_ || gen_ != seen; });
                if (stop_) return;
                seen = gen_;
                job = job_;
            }
            (*job)(id);
            {
                std::lock_guard<std::mutex> lk(m_);
                if (--pending_ == 0) done_.notify_one();
            }
        }
    }

public:
    // f(0) on the calling thread, f(1) .. f(n-1) on the workers; returns when all have finished
    void run(int n, const std::function<void(int)>& f)
    {
        while ((int)th_.size() < n - 1) { const int id = (int)th_.size() + 1; th_.emplace_back([this, id] { loop(id); }); }
        {
            std::lock_guard<std::mutex> lk(m_);
            job_ = &f; pending_ = n - 1; ++gen_;
        }
        cv_.notify_all();
        f(0);
        std::unique_lock<std::mutex> lk(m_);
        done_.wait(lk, [&] { return pending_ == 0; });
    }
    ~WorkerPool()
    {
        { std::lock_guard<std::mutex> lk(m_); stop_ = true; }
        cv_.notify_all();
        for (auto& t : th_) t.join();
    }
};

struct dwls_ctx {
    int device = 0;
    int num_sms = 148;
    cudaStream_t stream = nullptr;
    cudaStream_t stream2 = nullptr;                         // second lane of the pipelined host-batch entry point
    std::string err;
    int64_t launches = 0;
    int64_t h2d_bytes = 0, d2h_bytes = 0;                   // moved by the host-buffer entry points
    int group_frames = 0;
    bool interleave = false;                                // confidence path: the two solver planes in one sector-interleaved array (SolveArgs::il)
    float* lut = nullptr; float lut_sigma = -1.f;           // per-sigma LUT cache (fgs_filter.cpp:150-154)
    void* ws = nullptr; size_t ws_bytes = 0;                // solver workspace (grow-only)
    void* stage = nullptr; size_t stage_bytes = 0;          // device staging of the host entry points
    void* bmws = nullptr; size_t bmws_bytes = 0;            // stereo matcher: prefiltered views
    void* bmstage = nullptr; size_t bmstage_bytes = 0;      // stereo matcher: staging of its host entry point
    // pinned host staging of the single-call entry points: pageable caller memory is copied into it by a few host threads,
    // each feeding its own copy stream (the driver's own pageable path is single-threaded)
    void* hpin = nullptr; size_t hpin_bytes = 0;
    WorkerPool pool;
    cudaStream_t xs[8] = {};
    cudaEvent_t xe[8][2] = {};
    cudaEvent_t xready = nullptr;
    // optional per-kernel CUDA-event timing (bench.py's roofline block)
    bool timing = false;
    struct Span { int id; cudaEvent_t a, b; };
    std::vector<Span> spans;
    std::vector<cudaEvent_t> free_events;
    // CUDA-graph cache of whole filter_frames() launch sequences, keyed by every argument that reaches a kernel.
    // A sequence is captured the second time its key is seen (callers that pass fresh buffers every call never pay
    // for a capture) and replayed as one graph launch afterwards.
    bool graphs = true;
    cudaStream_t cap_stream = nullptr;
    struct GraphEntry { std::string key; cudaGraphExec_t exec; int64_t kernels; uint64_t last_use; };
    std::vector<GraphEntry> graph_cache;
    std::vector<size_t> seen_once;
    uint64_t graph_clock = 0;
    int64_t graph_replays = 0;
    // per-context (= per-device) kernel set-up: function attributes are per device, and contexts may live on different threads
    struct HCache { int ctas_per_sm = 0, K = 0; size_t smem = 0; };
    std::unordered_map<const void*, HCache> hcache;          // row-pass occupancy, by kernel instance
    std::unordered_set<const void*> smem_attr_done;          // kernels whose dynamic shared-memory limit has been raised
    // Calls share the workspace, the LUT and the cached graphs, so two calls on DIFFERENT streams must not overlap: every
    // device entry point records `busy` on its stream when it has queued its work, and the next one waits for it when it
    // runs on another stream (same stream: already ordered).
    cudaEvent_t busy = nullptr; cudaStream_t busy_stream = nullptr; bool busy_valid = false;
    // peer-memory gather: copy stream of the pushes, producer / completion events, bytes pushed
    cudaStream_t peer_stream = nullptr; cudaEvent_t peer_ev = nullptr, peer_done = nullptr; int64_t peer_bytes = 0;
    // device batch entry point: launch groups are dealt round-robin to `lanes` internal streams (forked from and joined
    // back into the caller's stream), each with its own workspace slice, so that the tail of one group's kernels overlaps
    // the head of another's; with small groups the planes a group hands from kernel to kernel stay in the 126 MB L2
    int lanes = 1;
    cudaStream_t lane_st[DWLS_MAX_LANES] = {};
    cudaEvent_t lane_ev[DWLS_MAX_LANES] = {};
    cudaEvent_t fork_ev = nullptr;
    // host-side launch set-up that does not change between calls: encoded tensor maps and occupancy answers
    struct TmapKey { const void* base; int pitch; int64_t rows; int bc, br; bool operator==(const TmapKey& o) const { return base == o.base && pitch == o.pitch && rows == o.rows && bc == o.bc && br == o.br; } };
    std::vector<std::pair<TmapKey, CUtensorMap>> tmaps;       // small LRU-less cache, cleared when the workspace moves
    struct VCache { int per_sm = 0; size_t smem = 0; int threads = 0; };
    std::unordered_map<const void*, VCache> vcache;
};

// orders a call on `st` after the previous call's work (see dwls_ctx::busy)
static cudaError_t call_enter(dwls_ctx* ctx, cudaStream_t st)
{
    if (ctx->busy_valid && ctx->busy_stream != st) return cudaStreamWaitEvent(st, ctx->busy, 0);
    return cudaSuccess;
}
static cudaError_t call_leave(dwls_ctx* ctx, cudaStream_t st)
{
    if (!ctx->busy) { cudaError_t e = cudaEventCreateWithFlags(&ctx->busy, cudaEventDisableTiming); if (e != cudaSuccess) return e; }
    ctx->busy_stream = st; ctx->busy_valid = true;
    return cudaEventRecord(ctx->busy, st);
}

enum KernelId { K_LUT = 0, K_WEIGHTS, K_CONF_R, K_CONF_L, K_UPSCALE, K_PREP, K_FILL, K_HSOLVE, K_VSOLVE, K_VSOLVE_FIN, K_UNPACK, K_PACK, K_TRANSPOSE,
                K_BM_PREF, K_BM, K_BM_FILL, K_HALF, K_COUNT };
static const char* const kKernelNames[K_COUNT] = { "lut_kernel", "weights_kernel", "conf_kernel<right>", "conf_kernel<left+lrc>", "upscale_prep_kernel",
                                                   "prep_noconf_kernel", "border_kernel", "hsolve_kernel", "vsolve_kernel", "vsolve_kernel<final>",
                                                   "unpack_kernel", "pack_kernel", "transpose_kernel", "bm_prefilter_kernel", "bm_kernel", "bm_fill_kernel",
                                                   "half_kernel" };

// counts the launch and, when timing is on, brackets it with CUDA events on the launching stream
struct KTimer {
    dwls_ctx* c; cudaStream_t st; cudaEvent_t stop = nullptr;
    static cudaEvent_t get(dwls_ctx* c) {
        if (!c->free_events.empty()) { cudaEvent_t e = c->free_events.back(); c->free_events.pop_back(); return e; }
        cudaEvent_t e = nullptr; cudaEventCreate(&e); return e;
    }
    KTimer(dwls_ctx* ctx, int id, cudaStream_t s) : c(ctx), st(s) {
        c->launches++;
        if (!c->timing) return;
        cudaEvent_t a = get(c); stop = get(c);
        cudaEventRecord(a, st);
        c->spans.push_back({ id, a, stop });
    }
    ~KTimer() { if (stop) cudaEventRecord(stop, st); }
};
#define TIMED(id, st, ...) do { KTimer kt_(ctx, id, st); __VA_ARGS__; } while (0)

static std::string g_create_err;


#define CK(call)                                                                                     \
    do {                                                                                             \
        cudaError_t e_ = (call);                                                                     \
        if (e_ != cudaSuccess) {                                                                     \
            char b_[512];                                                                            \
            snprintf(b_, sizeof b_, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
            ctx->err = b_;                                                                           \
            return DWLS_CUDA_ERROR;                                                                  \
        }                                                                                            \
    } while (0)

static dwls_status fail(dwls_ctx* ctx, dwls_status s, const char* msg) { if (ctx) ctx->err = msg; return s; }

static inline int round_up(int v, int m) { return (v + m - 1) / m * m; }
static inline size_t align256(size_t v) { return (v + 255) & ~(size_t)255; }

static dwls_status ensure(dwls_ctx* ctx, void** p, size_t* have, size_t need)
{
    if (*have >= need) return DWLS_OK;
    if (*p) {
        CK(cudaDeviceSynchronize());
        for (auto& ge : ctx->graph_cache) cudaGraphExecDestroy(ge.exec);      // cached graphs point into the old allocation
        ctx->graph_cache.clear();
        ctx->tmaps.clear();
        CK(cudaFree(*p)); *p = nullptr; *have = 0;
    }
    CK(cudaMalloc(p, need));
    *have = need;
    return DWLS_OK;
}

extern "C" {

void dwls_default_params(dwls_params* p)
{
    if (!p) return;
    p->lambda = 8000.0; p->sigma_color = 1.0; p->lrc_thresh = 24; p->disc_radius = 5;       // disparity_filters.cpp:154-155, 215
    p->left_offset = p->right_offset = p->top_offset = p->bottom_offset = 0;
    p->use_confidence = 1; p->lambda_attenuation = 0.25; p->num_iter = 3;                   // edge_filter.hpp:393
}

const char* dwls_version(void) { return "dwls_b200 0.1 (sm_100a)"; }

dwls_status dwls_ctx_create(int device, dwls_ctx** out)
{
    if (!out) return DWLS_BAD_ARG;
    *out = nullptr;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n <= 0) {
        g_create_err = std::string("no CUDA device: ") + cudaGetErrorString(e) + " — this library has no CPU fallback";
        return DWLS_CUDA_ERROR;
    }
    if (device < 0 || device >= n) { g_create_err = "device index out of range"; return DWLS_BAD_ARG; }
    dwls_ctx* ctx = new dwls_ctx();
    ctx->device = device;
    if ((e = cudaSetDevice(device)) != cudaSuccess || (e = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking)) != cudaSuccess) {
        g_create_err = std::string("cudaSetDevice/cudaStreamCreate: ") + cudaGetErrorString(e);
        delete ctx;
        return DWLS_CUDA_ERROR;
    }
    cudaDeviceGetAttribute(&ctx->num_sms, cudaDevAttrMultiProcessorCount, device);
    *out = ctx;
    return DWLS_OK;
}

void dwls_ctx_destroy(dwls_ctx* ctx)
{
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    if (ctx->stream) { cudaStreamSynchronize(ctx->stream); cudaStreamDestroy(ctx->stream); }
    if (ctx->stream2) { cudaStreamSynchronize(ctx->stream2); cudaStreamDestroy(ctx->stream2); }
    for (auto& ge : ctx->graph_cache) cudaGraphExecDestroy(ge.exec);
    if (ctx->cap_stream) cudaStreamDestroy(ctx->cap_stream);
    cudaFree(ctx->lut); cudaFree(ctx->ws); cudaFree(ctx->stage); cudaFree(ctx->bmws); cudaFree(ctx->bmstage);
    if (ctx->hpin) cudaFreeHost(ctx->hpin);
    for (int t = 0; t < 8; ++t) {
        if (ctx->xs[t]) cudaStreamDestroy(ctx->xs[t]);
        for (int k = 0; k < 2; ++k) if (ctx->xe[t][k]) cudaEventDestroy(ctx->xe[t][k]);
    }
    if (ctx->xready) cudaEventDestroy(ctx->xready);
    if (ctx->busy) cudaEventDestroy(ctx->busy);
    if (ctx->peer_stream) { cudaStreamSynchronize(ctx->peer_stream); cudaStreamDestroy(ctx->peer_stream); cudaEventDestroy(ctx->peer_ev); cudaEventDestroy(ctx->peer_done); }
    for (int l = 0; l < DWLS_MAX_LANES; ++l) { if (ctx->lane_st[l]) cudaStreamDestroy(ctx->lane_st[l]); if (ctx->lane_ev[l]) cudaEventDestroy(ctx->lane_ev[l]); }
    if (ctx->fork_ev) cudaEventDestroy(ctx->fork_ev);
    for (auto& sp : ctx->spans) { cudaEventDestroy(sp.a); cudaEventDestroy(sp.b); }
    for (auto e : ctx->free_events) cudaEventDestroy(e);
    delete ctx;
}

const char* dwls_last_error(const dwls_ctx* ctx) { return ctx ? ctx->err.c_str() : g_create_err.c_str(); }
int64_t dwls_kernel_launches(const dwls_ctx* ctx) { return ctx ? ctx->launches : 0; }
dwls_status dwls_set_group_frames(dwls_ctx* ctx, int frames)
{
    if (!ctx || frames < 0) return DWLS_BAD_ARG;
    ctx->group_frames = frames;
    return DWLS_OK;
}

dwls_status dwls_set_lanes(dwls_ctx* ctx, int lanes)
{
    if (!ctx || lanes < 1 || lanes > DWLS_MAX_LANES) return DWLS_BAD_ARG;
    ctx->lanes = lanes;
    return DWLS_OK;
}

dwls_status dwls_set_plane_interleave(dwls_ctx* ctx, int on)
{
    if (!ctx) return DWLS_BAD_ARG;
    ctx->interleave = on != 0;
    return DWLS_OK;
}

dwls_status dwls_set_cuda_graphs(dwls_ctx* ctx, int on)
{
    if (!ctx) return DWLS_BAD_ARG;
    ctx->graphs = on != 0;
    return DWLS_OK;
}
int64_t dwls_graph_replays(const dwls_ctx* ctx) { return ctx ? ctx->graph_replays : 0; }
void dwls_transfer_bytes(const dwls_ctx* ctx, int64_t* h2d, int64_t* d2h)
{
    if (h2d) *h2d = ctx ? ctx->h2d_bytes : 0;
    if (d2h) *d2h = ctx ? ctx->d2h_bytes : 0;
}

dwls_status dwls_set_kernel_timing(dwls_ctx* ctx, int on)
{
    if (!ctx) return DWLS_BAD_ARG;
    ctx->timing = on != 0;
    return DWLS_OK;
}

// synchronises the ctx streams, folds the recorded spans into per-kernel totals and clears them.
// names[i] / launches[i] / total_ms[i] for i < returned count (<= cap)
int dwls_kernel_timing_report(dwls_ctx* ctx, const char** names, int64_t* launches, double* total_ms, int cap)
{
    if (!ctx) return 0;
    cudaSetDevice(ctx->device);
    cudaDeviceSynchronize();
    int64_t n[K_COUNT] = { 0 }; double ms[K_COUNT] = { 0 };
    for (auto& sp : ctx->spans) {
        float t = 0.f;
        if (cudaEventElapsedTime(&t, sp.a, sp.b) == cudaSuccess) { n[sp.id]++; ms[sp.id] += t; }
        ctx->free_events.push_back(sp.a); ctx->free_events.push_back(sp.b);
    }
    ctx->spans.clear();
    int k = 0;
    for (int i = 0; i < K_COUNT && k < cap; ++i)
        if (n[i]) { names[k] = kKernelNames[i]; launches[k] = n[i]; total_ms[k] = ms[i]; ++k; }
    return k;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------
// shared pieces
// ---------------------------------------------------------------------------------------------
static const int LUT_LEVELS = 3 * 256 * 256;

static dwls_status ensure_lut(dwls_ctx* ctx, float sigma, cudaStream_t st)
{
    if (!ctx->lut) CK(cudaMalloc(&ctx->lut, sizeof(float) * LUT_LEVELS));
    if (ctx->lut_sigma != sigma) {
        TIMED(K_LUT, st, lut_kernel<<<(LUT_LEVELS + 255) / 256, 256, 0, st>>>(ctx->lut, LUT_LEVELS, sigma));
        CK(cudaGetLastError());
        ctx->lut_sigma = sigma;
    }
    return DWLS_OK;
}

// wstrip: columns per strip of the strip-major Cvert plane, 0 = row-major.
// vT: more than 2176 rows, the column pass runs as a row pass over transposed planes (tM elements per thread, tK threads,
// tpitch floats per transposed row)
struct SolvePlan { int hK, hM; int vM, vCols, vK; int wstrip; bool vT; int tK, tM, tpitch; };
constexpr int MAX_SIDE = 8192;

static bool plan_solver(int w, int h, SolvePlan* p)
{
    p->hM = w <= 2048 ? 8 : 16;
    p->hK = round_up((w + p->hM - 1) / p->hM, 32);
    if (p->hK > 512) return false;
    p->vT = false; p->tK = p->tM = p->tpitch = 0;
    if (h <= 9 * 128) { p->vM = 9; p->vCols = 8; }
    else if (h <= 17 * 128) { p->vM = 17; p->vCols = 4; }
    else if (h <= MAX_SIDE) {
        p->vT = true; p->vM = 0; p->vCols = 0; p->vK = 0; p->wstrip = 0;
        p->tM = h <= 2048 ? 8 : 16;
        p->tK = round_up((h + p->tM - 1) / p->tM, 32);
        p->tpitch = round_up(h, 8);
        return true;
    }
    else return false;
    p->vK = round_up((h + p->vM - 1) / p->vM, 32);
    p->wstrip = (p->vM == 9) ? 8 : 0;       // taller images mix 8- and 4-column strips over the same weights: keep them row-major
    return true;
}

template <int S, int MAXT, int HM, bool IL = false, bool DC = false>
static dwls_status launch_h_inst(dwls_ctx* ctx, const SolveArgs& a, const SolvePlan& pl, int frames, cudaStream_t st)
{
    const size_t smem = hsolve_smem_bytes<S, HM>(pl.hK, a.pitch);
    auto k = hsolve_kernel<S, MAXT, HM, IL, DC>;
    if (smem > 227 * 1024) return fail(ctx, DWLS_UNSUPPORTED, "row too wide for the shared-memory stages of the row pass");
    dwls_ctx::HCache& hc = ctx->hcache[(const void*)k];
    if (!hc.ctas_per_sm || hc.K != pl.hK || hc.smem != smem) {
        int per_sm = 0;
        CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k, pl.hK, smem));
        if (per_sm < 1) return fail(ctx, DWLS_CUDA_ERROR, "row pass does not fit on an SM");
        hc.ctas_per_sm = per_sm; hc.K = pl.hK; hc.smem = smem;
    }
    const int rows_total = a.h * frames;
    int grid = ctx->num_sms * hc.ctas_per_sm;
    if (grid > rows_total) grid = rows_total;
    KTimer kt_(ctx, K_HSOLVE, st);
    k<<<grid, pl.hK, smem, st>>>(a, rows_total);
    CK(cudaGetLastError());
    return DWLS_OK;
}

template <int S>
static dwls_status launch_h(dwls_ctx* ctx, const SolveArgs& a, const SolvePlan& pl, int frames, cudaStream_t st)
{
    if (a.il) {
        if constexpr (S == 2) { if (pl.hM == 8) return launch_h_inst<2, 256, 8, true>(ctx, a, pl, frames, st); }
        return fail(ctx, DWLS_CUDA_ERROR, "internal: interleaved planes need two planes and rows of at most 2048 pixels");
    }
    if (a.dl16) {
        if constexpr (S == 2) { if (pl.hM == 8) return launch_h_inst<2, 256, 8, false, true>(ctx, a, pl, frames, st); }
        return fail(ctx, DWLS_CUDA_ERROR, "internal: the d*c-forming row pass needs two planes and rows of at most 2048 pixels");
    }
    if (pl.hM == 8) return launch_h_inst<S, 256, 8>(ctx, a, pl, frames, st);          // w <= 2048
    if (pl.hK <= 256) return launch_h_inst<S, 256, 16>(ctx, a, pl, frames, st);       // w <= 4096
    return launch_h_inst<S, 512, 16>(ctx, a, pl, frames, st);                         // w <= 8192: level 2 takes the generic loop
}

// cuTensorMapEncodeTiled through the runtime's driver entry point (no link-time dependency on libcuda)
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn g_encode = nullptr;
static std::once_flag g_encode_once;
// resolved once per process (contexts may be created and used from several threads)
static EncodeTiledFn get_encode()
{
    std::call_once(g_encode_once, [] {
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q) == cudaSuccess && fn && q == cudaDriverEntryPointSuccess)
            g_encode = (EncodeTiledFn)fn;
    });
    return g_encode;
}

// 2-D fp32 tensor [rows][pitch] with a (box_rows x box_cols) tile
static dwls_status make_tmap_uncached(dwls_ctx* ctx, CUtensorMap* tm, const float* base, int pitch, int64_t rows, int box_cols, int box_rows);
static dwls_status make_tmap(dwls_ctx* ctx, CUtensorMap* tm, const float* base, int pitch, int64_t rows, int box_cols, int box_rows)
{
    const dwls_ctx::TmapKey key = { base, pitch, rows, box_cols, box_rows };
    for (auto& e : ctx->tmaps) if (e.first == key) { *tm = e.second; return DWLS_OK; }
    const dwls_status s = make_tmap_uncached(ctx, tm, base, pitch, rows, box_cols, box_rows);
    if (s) return s;
    if (ctx->tmaps.size() >= 256) ctx->tmaps.erase(ctx->tmaps.begin(), ctx->tmaps.begin() + 128);
    ctx->tmaps.push_back({ key, *tm });
    return DWLS_OK;
}
static dwls_status make_tmap_uncached(dwls_ctx* ctx, CUtensorMap* tm, const float* base, int pitch, int64_t rows, int box_cols, int box_rows)
{
    if (!get_encode()) return fail(ctx, DWLS_CUDA_ERROR, "cuTensorMapEncodeTiled is not available in this driver");
    const cuuint64_t dims[2] = { (cuuint64_t)pitch, (cuuint64_t)rows };
    const cuuint64_t strides[1] = { (cuuint64_t)pitch * sizeof(float) };
    const cuuint32_t box[2] = { (cuuint32_t)box_cols, (cuuint32_t)box_rows };
    const cuuint32_t estr[2] = { 1, 1 };
    const CUresult r = g_encode(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, (void*)base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        char b_[128];
        snprintf(b_, sizeof b_, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
        ctx->err = b_;
        return DWLS_CUDA_ERROR;
    }
    return DWLS_OK;
}

template <int M, int S, int VC, int FIN, int MAXT, bool WB, bool IL = false>
static dwls_status launch_v_inst(dwls_ctx* ctx, const SolveArgs& a, const SolvePlan& pl, int frames, cudaStream_t st)
{
    VTma g;
    g.nb = (a.h + 255) / 256;
    g.rb = round_up((a.h + g.nb - 1) / g.nb, 128 / (VC * 4));     // every TMA box must land on a 128-byte boundary in smem
    g.strips_x = (a.w + VC - 1) / VC;
    g.strips_total = g.strips_x * frames;
    const size_t smem = vsolve_smem_bytes<M, S, VC>(pl.vK, g.nb * g.rb, WB ? 2 : 1);
    if (smem > 227 * 1024) return fail(ctx, DWLS_UNSUPPORTED, "image too tall for the shared-memory stage of the column pass");
    auto k = vsolve_kernel<M, S, VC, FIN, MAXT, WB, IL>;
    dwls_ctx::VCache& vc = ctx->vcache[(const void*)k];
    if (!vc.per_sm || vc.smem != smem || vc.threads != VC * pl.vK) {
        int per_sm = 1;
        CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k, VC * pl.vK, smem));
        if (per_sm < 1) return fail(ctx, DWLS_CUDA_ERROR, "column pass does not fit on an SM");
        vc.per_sm = per_sm; vc.smem = smem; vc.threads = VC * pl.vK;
    }
    CUtensorMap tw, t0, t1;
    dwls_status s;
    const int64_t rows = (int64_t)frames * a.h;           // frames are contiguous planes: a.p_frame == a.w_frame == pitch * h
    if (WB) tw = CUtensorMap();
    else if ((s = make_tmap(ctx, &tw, a.weight, a.pitch, rows, VC, g.rb))) return s;
    if (IL) { if ((s = make_tmap(ctx, &t0, a.in[0], S * a.pitch, rows, S * VC, g.rb))) return s; t1 = t0; }   // one box row = the strip of all planes
    else {
        if ((s = make_tmap(ctx, &t0, a.in[0], a.pitch, rows, VC, g.rb))) return s;
        if (S > 1) { if ((s = make_tmap(ctx, &t1, a.in[1], a.pitch, rows, VC, g.rb))) return s; }
        else t1 = t0;
    }
    int grid = ctx->num_sms * vc.per_sm;
    if (grid > g.strips_total) grid = g.strips_total;
    TIMED(FIN == FIN_FLOAT ? K_VSOLVE : K_VSOLVE_FIN, st, k<<<grid, VC * pl.vK, smem, st>>>(a, g, tw, t0, t1));
    CK(cudaGetLastError());
    return DWLS_OK;
}

template <int S, int FIN>
static dwls_status launch_v(dwls_ctx* ctx, const SolveArgs& a, const SolvePlan& pl, int frames, cudaStream_t st)
{
    if (a.il) {
        if constexpr (S == 2) { if (pl.vM == 9 && pl.wstrip == 8) return launch_v_inst<9, 2, 8, FIN, 1024, true, true>(ctx, a, pl, frames, st); }
        return fail(ctx, DWLS_CUDA_ERROR, "internal: interleaved planes need two planes and at most 1152 rows");
    }
    if (pl.vM == 9) return pl.wstrip == 8 ? launch_v_inst<9, S, 8, FIN, 1024, true>(ctx, a, pl, frames, st)
                                          : launch_v_inst<9, S, 8, FIN, 1024, false>(ctx, a, pl, frames, st);
    // 1153..2176 rows, 17 rows per thread: a single plane still fits 8-column strips (32-byte rows) in 64 registers
    if (pl.vCols == 4 && S == 1) return launch_v_inst<17, 1, 8, FIN == FIN_WLS ? FIN_FLOAT : FIN, 1024, false>(ctx, a, pl, frames, st);
    if (pl.vCols == 4) return launch_v_inst<17, S, 4, FIN, 512, false>(ctx, a, pl, frames, st);
    return fail(ctx, DWLS_UNSUPPORTED, "column pass supports at most 2176 rows");
}

// scratch of the transposed column pass (plan.vT): Cvert transposed once per call, the S planes every column sweep
struct TPass { float* cvertT; float* pT[2]; size_t tplane; };

template <int S, int FIN>
static dwls_status launch_transpose(dwls_ctx* ctx, const TrArgs& t, int frames, cudaStream_t st)
{
    TIMED(K_TRANSPOSE, st, transpose_kernel<S, FIN><<<dim3((t.cols + 31) / 32, (t.rows + 31) / 32, frames), 256, 0, st>>>(t));
    CK(cudaGetLastError());
    return DWLS_OK;
}

// Cvert [h][pitch] -> [w][tpitch], once after the weights kernel
static dwls_status transpose_cvert(dwls_ctx* ctx, const float* cvert, size_t plane, int pitch, int w, int h, const SolvePlan& pl, const TPass& T,
                                   int frames, cudaStream_t st)
{
    TrArgs t;
    memset(&t, 0, sizeof t);
    t.in[0] = cvert; t.in_frame = (int64_t)plane; t.in_pitch = pitch;
    t.out[0] = T.cvertT; t.out_frame = (int64_t)T.tplane; t.out_pitch = pl.tpitch;
    t.rows = h; t.cols = w;
    return launch_transpose<1, FIN_FLOAT>(ctx, t, frames, st);
}

// one column sweep of a tall image: transpose the planes, solve their rows, transpose back (epilogue fused there)
static dwls_status column_sweep_transposed(dwls_ctx* ctx, const SolveArgs& a, int S, int fin, const SolvePlan& pl, const TPass& T, int frames,
                                           cudaStream_t st)
{
    TrArgs t;
    memset(&t, 0, sizeof t);
    for (int s = 0; s < S; ++s) { t.in[s] = a.in[s]; t.out[s] = T.pT[s]; }
    t.in_frame = a.p_frame; t.in_pitch = a.pitch; t.out_frame = (int64_t)T.tplane; t.out_pitch = pl.tpitch;
    t.rows = a.h; t.cols = a.w;
    dwls_status s = (S == 1) ? launch_transpose<1, FIN_FLOAT>(ctx, t, frames, st) : launch_transpose<2, FIN_FLOAT>(ctx, t, frames, st);
    if (s) return s;
    SolveArgs r = a;
    r.weight = T.cvertT; r.w_frame = (int64_t)T.tplane; r.p_frame = (int64_t)T.tplane; r.pitch = pl.tpitch; r.w = a.h; r.h = a.w;
    for (int k = 0; k < S; ++k) { r.in[k] = T.pT[k]; r.out[k] = T.pT[k]; }
    SolvePlan rp = pl;
    rp.hK = pl.tK; rp.hM = pl.tM;
    if ((s = (S == 1) ? launch_h<1>(ctx, r, rp, frames, st) : launch_h<2>(ctx, r, rp, frames, st))) return s;
    memset(&t, 0, sizeof t);
    for (int k = 0; k < S; ++k) { t.in[k] = T.pT[k]; t.out[k] = a.out[k]; }
    t.in_frame = (int64_t)T.tplane; t.in_pitch = pl.tpitch; t.out_frame = a.p_frame; t.out_pitch = a.pitch;
    t.rows = a.w; t.cols = a.h;
    t.out16 = a.out16; t.out16_step = a.out16_step; t.out16_frame = a.out16_frame;
    if (fin == FIN_WLS) return launch_transpose<2, FIN_WLS>(ctx, t, frames, st);
    if (fin == FIN_S16) return launch_transpose<1, FIN_S16>(ctx, t, frames, st);
    return (S == 1) ? launch_transpose<1, FIN_FLOAT>(ctx, t, frames, st) : launch_transpose<2, FIN_FLOAT>(ctx, t, frames, st);
}

// num_iter x (row pass, column pass, lambda *= attenuation) on S <= 2 planes sharing the weights
// (fgs_filter.cpp:207-212).  fin != FIN_FLOAT fuses the output conversion into the last column pass.
static dwls_status run_sweeps(dwls_ctx* ctx, SolveArgs a, const float* chor, const float* cvert, int S, int fin,
                              const SolvePlan& pl, int frames, double lambda, double att, int num_iter, cudaStream_t st,
                              const TPass* T = nullptr)
{
    float lam = (float)lambda;                    // the reference keeps lambda and the attenuation in float (70-71, 146-147)
    const float fatt = (float)att;
    for (int n = 0; n < num_iter; ++n) {
        a.lam = lam;
        a.weight = chor;
        dwls_status s = (S == 1) ? launch_h<1>(ctx, a, pl, frames, st) : launch_h<2>(ctx, a, pl, frames, st);
        if (s) return s;
        for (int k = 0; k < S; ++k) { a.in[k] = a.out[k]; a.in_frame[k] = 0; a.in_pitch[k] = 0; }   // only the first row pass may read from elsewhere
        a.dl16 = nullptr;
        a.weight = cvert;
        const bool last = (n == num_iter - 1);
        if (pl.vT) {
            if (!T) return fail(ctx, DWLS_CUDA_ERROR, "internal: transposed column pass without scratch planes");
            s = column_sweep_transposed(ctx, a, S, last ? fin : FIN_FLOAT, pl, *T, frames, st);
        }
        else if (last && fin == FIN_WLS) s = launch_v<2, FIN_WLS>(ctx, a, pl, frames, st);
        else if (last && fin == FIN_S16) s = launch_v<1, FIN_S16>(ctx, a, pl, frames, st);
        else s = (S == 1) ? launch_v<1, FIN_FLOAT>(ctx, a, pl, frames, st) : launch_v<2, FIN_FLOAT>(ctx, a, pl, frames, st);
        if (s) return s;
        lam *= fatt;
    }
    return DWLS_OK;
}

static dwls_status launch_weights(dwls_ctx* ctx, const uint8_t* guide, int64_t gstep, int64_t gframe, int cn, int w, int h,
                                  int pitch, float* chor, float* cvert, int64_t wframe, int frames, cudaStream_t st, int vstrip)
{
    if (vstrip <= 0) vstrip = pitch;
    // threads per CTA: spread the row over equally wide tiles of at most 1024 pixels.  (Narrower tiles for launches too small
    // to fill the machine were measured in round 2 and lose: one 1080p frame 12.9 -> 14.8 us, the conf kernels 33 -> 59 us.)
    const int tiles = (pitch + 1023) / 1024;
    const int nt = round_up((pitch + 4 * tiles - 1) / (4 * tiles), 32);
    const int rowpitch = weights_row_pitch(nt, cn);
    const size_t smem = (size_t)(WT_RB + 1) * rowpitch + sizeof(float) * WT_LUT + 8;
    dim3 grid((pitch + 4 * nt - 1) / (4 * nt), (h + WT_RB - 1) / WT_RB, frames);
    KTimer kt_(ctx, K_WEIGHTS, st);
    if (cn == 1) weights_kernel<1><<<grid, nt, smem, st>>>(guide, gstep, gframe, w, h, pitch, rowpitch, ctx->lut, chor, cvert, wframe, vstrip);
    else weights_kernel<3><<<grid, nt, smem, st>>>(guide, gstep, gframe, w, h, pitch, rowpitch, ctx->lut, chor, cvert, wframe, vstrip);
    CK(cudaGetLastError());
    return DWLS_OK;
}

static dwls_status launch_border(dwls_ctx* ctx, const BorderArgs& ba, int frames, cudaStream_t st)
{
    const int tasks = (ba.H - ba.rh) + (ba.W > ba.rw ? ba.rh : 0);       // rows of the bands, one warp each
    if (tasks <= 0) return DWLS_OK;
    int blocks = (tasks + 7) / 8;
    const int cap = (ctx->num_sms * 8 + frames - 1) / frames;
    if (blocks > cap) blocks = cap < 1 ? 1 : cap;
    TIMED(K_FILL, st, border_kernel<<<dim3(blocks, frames), 256, 0, st>>>(ba));
    CK(cudaGetLastError());
    return DWLS_OK;
}

// int16 disparity maps of a launch group as one 3-D tensor {cols, rows, frames} with a {256 x box_rows x 1} box;
// returns false when the layout does not meet the TMA alignment rules (the caller then uses the plain-load path)
static bool make_tmap_s16(dwls_ctx* ctx, CUtensorMap* tm, const int16_t* base, int cols, int rows, int frames, int64_t step_elems,
                          int64_t frame_elems, int box_rows)
{
    if (((uintptr_t)base & 15) || (step_elems & 7) || step_elems <= 0) return false;
    if (frames > 1 && ((frame_elems & 7) || frame_elems <= 0)) return false;
    if (!get_encode()) return false;
    const cuuint64_t dims[3] = { (cuuint64_t)cols, (cuuint64_t)rows, (cuuint64_t)frames };
    const cuuint64_t strides[2] = { (cuuint64_t)step_elems * 2, (cuuint64_t)(frames > 1 ? frame_elems : step_elems * rows) * 2 };
    const cuuint32_t box[3] = { 256, (cuuint32_t)box_rows, 1 };
    const cuuint32_t estr[3] = { 1, 1, 1 };
    return g_encode(tm, CU_TENSOR_MAP_DATA_TYPE_UINT16, 3, (void*)base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                    CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// confidence stage: right view first (its map is gathered by the left view's LRC step)
template <int R, bool LEFT, bool TMA>
static void launch_conf_inst(dwls_ctx* ctx, const ConfArgs& a, int G, dim3 grid, cudaStream_t st, const CUtensorMap& tm)
{
    const size_t smem = conf_run_smem(R, LEFT);
    auto k = conf_run_kernel<R, LEFT, TMA>;
    if (ctx->smem_attr_done.insert((const void*)k).second) cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    k<<<grid, CR_T, smem, st>>>(a, G, tm);
}

template <int R>
static void launch_conf_run(dwls_ctx* ctx, const ConfArgs& a, int G, dim3 grid, cudaStream_t st, bool left, const CUtensorMap* tm)
{
    CUtensorMap none;
    memset(&none, 0, sizeof none);
    if (left) { if (tm) launch_conf_inst<R, true, true>(ctx, a, G, grid, st, *tm); else launch_conf_inst<R, true, false>(ctx, a, G, grid, st, none); }
    else { if (tm) launch_conf_inst<R, false, true>(ctx, a, G, grid, st, *tm); else launch_conf_inst<R, false, false>(ctx, a, G, grid, st, none); }
}

static dwls_status launch_conf(dwls_ctx* ctx, const ConfArgs& a, int frames, cudaStream_t st)
{
    if (a.radius <= CR_MAXR) {
        const int runs = (a.rw + 7) / 8, gmax = CR_GMAX;
        const int tiles = (runs + gmax - 1) / gmax, G = (runs + tiles - 1) / tiles;
        const dim3 grid(tiles, (a.rh + CR_TH - 1) / CR_TH, frames);
        const int box_rows = CR_TH + 2 * a.radius + 1;
        for (int left = 0; left < 2; ++left) {
            CUtensorMap tm;
            const bool tma = left ? make_tmap_s16(ctx, &tm, a.dl, a.dw, a.dh, frames, a.dl_step, a.dl_frame, box_rows)
                                  : make_tmap_s16(ctx, &tm, a.dr, a.dw, a.dh, frames, a.dr_step, a.dr_frame, box_rows);
            const CUtensorMap* ptm = tma ? &tm : nullptr;
            KTimer kt_(ctx, left ? K_CONF_L : K_CONF_R, st);
            switch (a.radius) {
            case 0: launch_conf_run<0>(ctx, a, G, grid, st, left, ptm); break;
            case 1: launch_conf_run<1>(ctx, a, G, grid, st, left, ptm); break;
            case 2: launch_conf_run<2>(ctx, a, G, grid, st, left, ptm); break;
            case 3: launch_conf_run<3>(ctx, a, G, grid, st, left, ptm); break;
            case 4: launch_conf_run<4>(ctx, a, G, grid, st, left, ptm); break;
            case 5: launch_conf_run<5>(ctx, a, G, grid, st, left, ptm); break;
            case 6: launch_conf_run<6>(ctx, a, G, grid, st, left, ptm); break;
            default: launch_conf_run<7>(ctx, a, G, grid, st, left, ptm); break;
            }
            CK(cudaGetLastError());
        }
        return DWLS_OK;
    }
    const dim3 cgrid((a.rw + CF_TW - 1) / CF_TW, (a.rh + CF_ROWS - 1) / CF_ROWS, frames);
    const size_t csmem = conf_smem_bytes(a.radius);
    if (csmem > 48 * 1024) {
        CK(cudaFuncSetAttribute(conf_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)csmem));
        CK(cudaFuncSetAttribute(conf_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)csmem));
    }
    TIMED(K_CONF_R, st, conf_kernel<false><<<cgrid, CF_T, csmem, st>>>(a));
    CK(cudaGetLastError());
    TIMED(K_CONF_L, st, conf_kernel<true><<<cgrid, CF_T, csmem, st>>>(a));
    CK(cudaGetLastError());
    return DWLS_OK;
}

// ---------------------------------------------------------------------------------------------
// DisparityWLSFilter::filter on device buffers
// ---------------------------------------------------------------------------------------------
struct Geom {
    int dw, dh, W, H, vcn;
    bool same;
    float resize_factor;
    int rx, ry, rw, rh, rrx;        // disparity-resolution ROI and its mirror
    int fx, fy, fw, fh;             // ROI at view resolution
    int pitch;                      // solver plane pitch (floats)
    SolvePlan plan;
};

// ROI rules of disparity_filters.cpp:224-233, 202-203, 245-246; returns BAD_SIZE where cv::Mat(m, roi) would assert
static dwls_status make_geom(dwls_ctx* ctx, int dw, int dh, int W, int H, int vcn, const int* roi_in, const dwls_params* p, Geom* g)
{
    g->dw = dw; g->dh = dh; g->W = W; g->H = H; g->vcn = vcn;
    g->same = (dw == W && dh == H);
    g->resize_factor = g->same ? 1.0f : dw / (float)W;
    if (roi_in && roi_in[2] * roi_in[3] != 0) { g->rx = roi_in[0]; g->ry = roi_in[1]; g->rw = roi_in[2]; g->rh = roi_in[3]; }
    else { g->rx = p->left_offset; g->ry = p->top_offset; g->rw = dw - p->left_offset - p->right_offset; g->rh = dh - p->top_offset - p->bottom_offset; }
    if (g->rx < 0 || g->ry < 0 || g->rw <= 0 || g->rh <= 0 || g->rx + g->rw > dw || g->ry + g->rh > dh)
        return fail(ctx, DWLS_BAD_SIZE, "ROI does not fit inside the disparity map");
    g->rrx = dw - (g->rx + g->rw);
    if (g->same) { g->fx = g->rx; g->fy = g->ry; g->fw = g->rw; g->fh = g->rh; }
    else {
        const float x_ratio = W / (float)dw, y_ratio = H / (float)dh;
        g->fx = (int)(g->rx * x_ratio); g->fy = (int)(g->ry * y_ratio); g->fw = (int)(g->rw * x_ratio); g->fh = (int)(g->rh * y_ratio);
        if (g->fx < 0 || g->fy < 0 || g->fw <= 0 || g->fh <= 0 || g->fx + g->fw > W || g->fy + g->fh > H)
            return fail(ctx, DWLS_BAD_SIZE, "upscaled ROI does not fit inside the view");
    }
    g->pitch = round_up(g->fw, 8);
    if (!plan_solver(g->fw, g->fh, &g->plan)) return fail(ctx, DWLS_UNSUPPORTED, "ROI larger than 8192 x 8192 is not supported by the on-chip solver");
    return DWLS_OK;
}

static dwls_status check_params(dwls_ctx* ctx, const dwls_params* p)
{
    if (!p) return fail(ctx, DWLS_BAD_ARG, "params is NULL");
    if (!(p->lambda >= 0) || !(p->sigma_color >= 0) || p->num_iter < 1)                       // fgs_filter.cpp:143
        return fail(ctx, DWLS_BAD_ARG, "lambda >= 0, sigma_color >= 0 and num_iter >= 1 are required");
    if (p->use_confidence && (p->disc_radius < 0 || p->disc_radius > 64)) return fail(ctx, DWLS_UNSUPPORTED, "depth discontinuity radius must be in [0, 64]");
    return DWLS_OK;
}

// DWLS_CONF_ALIAS=0 (A/B runs): always write solver plane c, even when the exported confidence map could stand in for it
static bool conf_alias_enabled()
{
    static const bool on = [] { const char* e = getenv("DWLS_CONF_ALIAS"); return !(e && e[0] == '0'); }();
    return on;
}

// per-frame workspace layout (floats): chor | cvert | plane0 | plane1 | conf_r | conf_lo
struct WsLayout { size_t plane, confr, conflo, tplane, per_frame; int cr_pitch; };   // tplane: transposed plane (tall ROIs), x3: Cvert + 2 planes
static WsLayout ws_layout(const Geom& g, bool use_conf)
{
    WsLayout l;
    l.plane = (size_t)g.pitch * g.fh;
    l.cr_pitch = round_up(g.rw, 8);
    l.confr = use_conf ? (size_t)l.cr_pitch * g.rh : 0;
    l.conflo = (use_conf && !g.same) ? (size_t)round_up(g.dw * g.dh, 8) : 0;   // keeps what follows 32-byte aligned
    l.tplane = g.plan.vT ? (size_t)g.plan.tpitch * g.fw : 0;
    l.per_frame = 4 * l.plane + l.confr + l.conflo + 3 * l.tplane;
    return l;
}

static dwls_status filter_frames(dwls_ctx* ctx, const dwls_batch* b, int f0, int nf, const Geom& g, const dwls_params* p,
                                 uint8_t* mask, int64_t mask_step, bool mask_only, cudaStream_t st, float* ws_base = nullptr)
{
    const bool conf = p->use_confidence != 0;
    const WsLayout L = ws_layout(g, conf);
    float* base = ws_base ? ws_base : (float*)ctx->ws;
    float* chor = base; float* cvert = chor + (size_t)nf * L.plane;
    float* p0 = cvert + (size_t)nf * L.plane; float* p1 = p0 + (size_t)nf * L.plane;
    float* confr = p1 + (size_t)nf * L.plane; float* conflo = confr + (size_t)nf * L.confr;
    TPass T;
    T.cvertT = conflo + (size_t)nf * L.conflo; T.pT[0] = T.cvertT + (size_t)nf * L.tplane; T.pT[1] = T.pT[0] + (size_t)nf * L.tplane;
    T.tplane = L.tplane;

    const int16_t* dl = (const int16_t*)((const uint8_t*)b->disp_left + (size_t)f0 * b->dl_frame);
    const int16_t* dr = b->disp_right ? (const int16_t*)((const uint8_t*)b->disp_right + (size_t)f0 * b->dr_frame) : nullptr;
    const uint8_t* view = b->view ? b->view + (size_t)f0 * b->view_frame : nullptr;
    int16_t* out = b->out ? (int16_t*)((uint8_t*)b->out + (size_t)f0 * b->out_frame) : nullptr;
    float* cexp = b->conf ? (float*)((uint8_t*)b->conf + (size_t)f0 * b->conf_frame) : nullptr;

    // confidence path, rows of at most 2048 pixels and at most 1152 of them: the two planes share one interleaved array (p0 | p1)
    const bool il = conf && !mask_only && ctx->interleave && g.plan.hM == 8 && g.plan.vM == 9 && !g.plan.vT && g.plan.wstrip == 8;
    bool out_border_done = false;
    // Full resolution with the confidence map exported: the map holds, inside the ROI, exactly the values of solver plane c, so
    // the confidence kernel need not write that plane a second time - the first row pass fetches its rows from the map
    // (7.6 MB per 1080p frame less DRAM traffic).  Needs 16-byte aligned ROI rows in the map and `pitch` readable floats per row.
    const int64_t cstep_e = b->conf_step / 4, cframe_e = b->conf_frame / 4;
    const bool alias_c = conf && !mask_only && !il && cexp && conf_alias_enabled() &&
                         (((uintptr_t)(cexp + (size_t)g.fy * cstep_e + g.fx)) & 15) == 0 && (cstep_e & 3) == 0 && (nf == 1 || (cframe_e & 3) == 0) &&
                         g.fx + g.pitch <= g.W && cstep_e < ((int64_t)1 << 30);      // (full resolution: fx, fy, W are rx, ry, dw)
    // ... and with 16-byte aligned ROI rows in the left disparity map (and rows of at most 2048 pixels) the first row pass also
    // forms plane d*c = c * float(L) itself instead of reading what the confidence kernel would have written (another 11.4 MB)
    const int64_t dstep_e = b->dl_step / 2, dframe_e = b->dl_frame / 2;
    const bool form_dc = alias_c && g.same && g.plan.hM == 8 && !g.plan.vT && (((uintptr_t)(dl + (size_t)g.ry * dstep_e + g.rx)) & 15) == 0 && (dstep_e & 7) == 0 &&
                         (nf == 1 || (dframe_e & 7) == 0);
    if (conf) {
        dwls_status s2;
        ConfArgs a;
        memset(&a, 0, sizeof a);
        a.dl = dl; a.dl_step = b->dl_step / 2; a.dl_frame = b->dl_frame / 2;
        a.dr = dr; a.dr_step = b->dr_step / 2; a.dr_frame = b->dr_frame / 2;
        a.dw = g.dw; a.dh = g.dh; a.rx = g.rx; a.ry = g.ry; a.rw = g.rw; a.rh = g.rh; a.rrx = g.rrx;
        a.radius = p->disc_radius;
        const int k = 2 * p->disc_radius + 1;
        a.scale_d = 1.0 / (k * k); a.scale_f = (float)a.scale_d;
        a.roll_off = 0.001f / (g.resize_factor * g.resize_factor);                             // 156, 359
        a.thresh = (int)(g.resize_factor * p->lrc_thresh);                                     // 318
        a.conf_r = confr; a.cr_frame = L.confr; a.cr_pitch = L.cr_pitch;
        if (g.same) { a.conf = cexp; a.conf_step = b->conf_step / 4; a.conf_frame = b->conf_frame / 4; a.plane_dc = p0; a.plane_c = p1; }
        else { a.conf = conflo; a.conf_step = g.dw; a.conf_frame = L.conflo; a.plane_dc = nullptr; a.plane_c = nullptr; }
        if (mask_only) { a.conf = nullptr; a.plane_dc = nullptr; a.plane_c = nullptr; }
        a.p_frame = L.plane; a.pitch = g.pitch; a.il = il;
        a.mask = mask; a.mask_step = mask_step;
        a.skip_c = (alias_c && g.same) ? 1 : 0; a.skip_dc = form_dc ? 1 : 0;
        if ((s2 = launch_conf(ctx, a, nf, st))) return s2;
        if (a.conf || a.mask) {          // zero the exported map (and the mask) outside the ROI (disparity_filters.cpp:187)
            BorderArgs ba;
            memset(&ba, 0, sizeof ba);
            ba.conf = a.conf; ba.conf_step = a.conf_step; ba.conf_frame = a.conf_frame; ba.mask = a.mask; ba.mask_step = a.mask_step;
            ba.W = g.dw; ba.H = g.dh; ba.rx = g.rx; ba.ry = g.ry; ba.rw = g.rw; ba.rh = g.rh;
            if (g.same && !mask_only) { ba.out = out; ba.out_step = b->out_step / 2; ba.out_frame = b->out_frame / 2; out_border_done = true; }
            dwls_status bs = launch_border(ctx, ba, nf, st);
            if (bs) return bs;
        }
        if (mask_only) return DWLS_OK;
    }
    if (!g.same) {
        UpArgs u;
        memset(&u, 0, sizeof u);
        u.disp = dl; u.d_step = b->dl_step / 2; u.d_frame = b->dl_frame / 2;
        u.conf_lo = conf ? conflo : nullptr; u.cl_step = g.dw; u.cl_frame = L.conflo;
        u.dw = g.dw; u.dh = g.dh; u.W = g.W; u.H = g.H;
        u.scale_x = 1.0 / ((double)g.W / g.dw); u.scale_y = 1.0 / ((double)g.H / g.dh);
        u.x_ratio = g.W / (float)g.dw;
        u.fx = g.fx; u.fy = g.fy; u.fw = g.fw; u.fh = g.fh;
        u.conf = conf ? cexp : nullptr; u.conf_step = b->conf_step / 4; u.conf_frame = b->conf_frame / 4;
        u.plane_dc = p0; u.plane_c = conf ? p1 : nullptr; u.p_frame = L.plane; u.pitch = g.pitch; u.il = il; u.skip_c = alias_c ? 1 : 0;
        TIMED(K_UPSCALE, st, upscale_prep_kernel<<<dim3((upscale_groups(g.W, g.fx) + 127) / 128, g.H, nf), 128, 0, st>>>(u));
        CK(cudaGetLastError());
    } else if (!conf) {
        TIMED(K_PREP, st, prep_noconf_kernel<<<dim3((g.fw + 255) / 256, g.fh, nf), 256, 0, st>>>(dl + (size_t)g.fy * (b->dl_step / 2) + g.fx, b->dl_step / 2,
                                                                            b->dl_frame / 2, g.fw, g.fh, p0, L.plane, g.pitch));
        CK(cudaGetLastError());
    }
    if (!out_border_done) {
        BorderArgs ba;
        memset(&ba, 0, sizeof ba);
        ba.out = out; ba.out_step = b->out_step / 2; ba.out_frame = b->out_frame / 2;
        ba.W = g.W; ba.H = g.H; ba.rx = g.fx; ba.ry = g.fy; ba.rw = g.fw; ba.rh = g.fh;
        dwls_status bs = launch_border(ctx, ba, nf, st);
        if (bs) return bs;
    }
    dwls_status s = launch_weights(ctx, view + (size_t)g.fy * b->view_step + (size_t)g.fx * g.vcn, b->view_step, b->view_frame, g.vcn,
                                   g.fw, g.fh, g.pitch, chor, cvert, L.plane, nf, st, g.plan.wstrip);
    if (s) return s;
    if (g.plan.vT && (s = transpose_cvert(ctx, cvert, L.plane, g.pitch, g.fw, g.fh, g.plan, T, nf, st))) return s;
    SolveArgs a;
    memset(&a, 0, sizeof a);
    a.w_frame = L.plane; a.p_frame = L.plane; a.pitch = g.pitch; a.w = g.fw; a.h = g.fh;
    a.in[0] = p0; a.out[0] = p0; a.in[1] = p1; a.out[1] = p1; a.il = il;
    if (alias_c) { a.in[1] = cexp + (size_t)g.fy * cstep_e + g.fx; a.in_frame[1] = cframe_e; a.in_pitch[1] = (int)cstep_e; }
    if (form_dc) { a.dl16 = dl + (size_t)g.ry * dstep_e + g.rx; a.dl16_frame = dframe_e; a.dl16_step = dstep_e; }
    a.out16 = out + (size_t)g.fy * (b->out_step / 2) + g.fx; a.out16_step = b->out_step / 2; a.out16_frame = b->out_frame / 2;
    return run_sweeps(ctx, a, chor, cvert, conf ? 2 : 1, conf ? FIN_WLS : FIN_S16, g.plan, nf, p->lambda, p->lambda_attenuation, p->num_iter, st, &T);
}

// filter_frames() through the graph cache: same launches, one cudaGraphLaunch once the sequence has been seen twice.
// Workspace and LUT must already be in place (no allocation happens inside a capture).
static dwls_status filter_frames_graphed(dwls_ctx* ctx, const dwls_batch* b, int f0, int nf, const Geom& g, const dwls_params* p,
                                         const int* roi_in, cudaStream_t st, float* ws_base = nullptr)
{
    if (!ctx->graphs || ctx->timing) return filter_frames(ctx, b, f0, nf, g, p, nullptr, 0, false, st, ws_base);
    std::string key;
    auto put = [&key](const void* v, size_t n) { key.append((const char*)v, n); };
#define PUT(x) put(&(x), sizeof(x))
    PUT(b->disp_left); PUT(b->dl_step); PUT(b->dl_frame); PUT(b->disp_right); PUT(b->dr_step); PUT(b->dr_frame);
    PUT(b->disp_cols); PUT(b->disp_rows); PUT(b->view); PUT(b->view_step); PUT(b->view_frame); PUT(b->view_cols); PUT(b->view_rows);
    PUT(b->view_channels); PUT(b->out); PUT(b->out_step); PUT(b->out_frame); PUT(b->conf); PUT(b->conf_step); PUT(b->conf_frame);
    PUT(f0); PUT(nf);
    PUT(p->lambda); PUT(p->sigma_color); PUT(p->lrc_thresh); PUT(p->disc_radius); PUT(p->left_offset); PUT(p->right_offset);
    PUT(p->top_offset); PUT(p->bottom_offset); PUT(p->use_confidence); PUT(p->lambda_attenuation); PUT(p->num_iter);
    const int no_roi[4] = { -1, -1, -1, -1 };
    put(roi_in ? roi_in : no_roi, sizeof no_roi);
    float* base = ws_base ? ws_base : (float*)ctx->ws;
    PUT(base); PUT(ctx->lut); PUT(ctx->interleave);
#undef PUT
    for (auto& ge : ctx->graph_cache)
        if (ge.key == key) {
            ge.last_use = ++ctx->graph_clock;
            CK(cudaGraphLaunch(ge.exec, st));
            ctx->launches += ge.kernels;
            ctx->graph_replays++;
            return DWLS_OK;
        }
    const size_t h = std::hash<std::string>()(key);
    bool seen = false;
    for (size_t v : ctx->seen_once) seen |= (v == h);
    if (!seen) {
        if (ctx->seen_once.size() >= 64) ctx->seen_once.erase(ctx->seen_once.begin());
        ctx->seen_once.push_back(h);
        return filter_frames(ctx, b, f0, nf, g, p, nullptr, 0, false, st, ws_base);
    }
    // capture on a private stream (the caller's may be the legacy default stream, which cannot be captured)
    if (!ctx->cap_stream) CK(cudaStreamCreateWithFlags(&ctx->cap_stream, cudaStreamNonBlocking));
    const int64_t before = ctx->launches;
    CK(cudaStreamBeginCapture(ctx->cap_stream, cudaStreamCaptureModeThreadLocal));
    const dwls_status s = filter_frames(ctx, b, f0, nf, g, p, nullptr, 0, false, ctx->cap_stream, ws_base);
    cudaGraph_t graph = nullptr;
    const cudaError_t ce = cudaStreamEndCapture(ctx->cap_stream, &graph);
    const int64_t kernels = ctx->launches - before;
    ctx->launches = before;
    if (s) { if (graph) cudaGraphDestroy(graph); return s; }
    if (ce != cudaSuccess || !graph) {                      // capture refused: run the plain launches instead
        cudaGetLastError();
        if (graph) cudaGraphDestroy(graph);
        return filter_frames(ctx, b, f0, nf, g, p, nullptr, 0, false, st, ws_base);
    }
    cudaGraphExec_t exec = nullptr;
    const cudaError_t ie = cudaGraphInstantiate(&exec, graph, 0);
    cudaGraphDestroy(graph);
    if (ie != cudaSuccess) { cudaGetLastError(); return filter_frames(ctx, b, f0, nf, g, p, nullptr, 0, false, st, ws_base); }
    if (ctx->graph_cache.size() >= 16) {
        size_t lru = 0;
        for (size_t i = 1; i < ctx->graph_cache.size(); ++i) if (ctx->graph_cache[i].last_use < ctx->graph_cache[lru].last_use) lru = i;
        // the evicted graph may still be running: the runtime defers the release until it has finished
        cudaGraphExecDestroy(ctx->graph_cache[lru].exec);
        ctx->graph_cache.erase(ctx->graph_cache.begin() + lru);
    }
    ctx->graph_cache.push_back({ key, exec, kernels, ++ctx->graph_clock });
    CK(cudaGraphLaunch(exec, st));
    ctx->launches += kernels;
    ctx->graph_replays++;
    return DWLS_OK;
}

static int pick_group(const dwls_ctx* ctx, size_t per_frame_bytes, int n)
{
    if (ctx->group_frames > 0) return ctx->group_frames < n ? ctx->group_frames : n;
    // default: keep a group's workspace under 2 GiB so launches are large but allocation stays modest,
    // and split the batch into groups of EQUAL size (a short trailing group is a launch too small to fill the GPU)
    size_t g = ((size_t)2 << 30) / (per_frame_bytes ? per_frame_bytes : 1);
    if (g < 1) g = 1;
    if (g > 48) g = 48;
    if (g >= (size_t)n) return n;
    const int groups = (n + (int)g - 1) / (int)g;
    return (n + groups - 1) / groups;
}

static dwls_status validate_batch(dwls_ctx* ctx, const dwls_batch* b, const dwls_params* p)
{
    if (!b || b->n_frames <= 0) return fail(ctx, DWLS_BAD_ARG, "empty batch");
    if (!b->disp_left || b->disp_cols <= 0 || b->disp_rows <= 0) return fail(ctx, DWLS_BAD_ARG, "left disparity map is empty");             // 221
    if (!b->view || b->view_cols <= 0 || b->view_rows <= 0 || (b->view_channels != 1 && b->view_channels != 3))
        return fail(ctx, DWLS_BAD_ARG, "left view must be a non-empty CV_8UC1 or CV_8UC3 image");                                            // 222
    if (p->use_confidence && !b->disp_right) return fail(ctx, DWLS_BAD_ARG, "filtering with confidence requires the right disparity map"); // 262
    if (!b->out) return fail(ctx, DWLS_BAD_ARG, "output is NULL");
    if ((b->dl_step | b->dl_frame | b->dr_step | b->dr_frame | b->out_step | b->out_frame) & 1) return fail(ctx, DWLS_BAD_ARG, "CV_16S strides must be even");
    if ((b->conf_step | b->conf_frame) & 3) return fail(ctx, DWLS_BAD_ARG, "CV_32F strides must be multiples of 4");
    return DWLS_OK;
}

extern "C" dwls_status dwls_filter_batch_dev(dwls_ctx* ctx, const dwls_batch* b, const int* roi_in, const dwls_params* p,
                                             int* roi_out, void* stream)
{
    if (!ctx) return DWLS_BAD_ARG;
    dwls_status s = check_params(ctx, p);
    if (s) return s;
    if ((s = validate_batch(ctx, b, p))) return s;
    Geom g;
    if ((s = make_geom(ctx, b->disp_cols, b->disp_rows, b->view_cols, b->view_rows, b->view_channels, roi_in, p, &g))) {
        if (roi_out) { roi_out[0] = g.rx; roi_out[1] = g.ry; roi_out[2] = g.rw; roi_out[3] = g.rh; }
        return s;
    }
    if (roi_out) { roi_out[0] = g.rx; roi_out[1] = g.ry; roi_out[2] = g.rw; roi_out[3] = g.rh; }
    CK(cudaSetDevice(ctx->device));
    cudaStream_t st = stream ? (cudaStream_t)stream : ctx->stream;
    const WsLayout L = ws_layout(g, p->use_confidence != 0);
    const int G = pick_group(ctx, L.per_frame * sizeof(float), b->n_frames);
    const int groups = (b->n_frames + G - 1) / G;
    const int lanes = ctx->timing ? 1 : std::min(ctx->lanes, groups);     // per-kernel timing wants the kernels one after another
    const size_t lane_ws = align256((size_t)G * L.per_frame * sizeof(float));
    if ((s = ensure(ctx, &ctx->ws, &ctx->ws_bytes, lanes * lane_ws))) return s;
    CK(call_enter(ctx, st));
    if ((s = ensure_lut(ctx, (float)p->sigma_color, st))) return s;
    if (lanes > 1) {
        if (!ctx->fork_ev) CK(cudaEventCreateWithFlags(&ctx->fork_ev, cudaEventDisableTiming));
        CK(cudaEventRecord(ctx->fork_ev, st));
        for (int l = 0; l < lanes; ++l) {
            if (!ctx->lane_st[l]) {
                CK(cudaStreamCreateWithFlags(&ctx->lane_st[l], cudaStreamNonBlocking));
                CK(cudaEventCreateWithFlags(&ctx->lane_ev[l], cudaEventDisableTiming));
            }
            CK(cudaStreamWaitEvent(ctx->lane_st[l], ctx->fork_ev, 0));
        }
    }
    int gi = 0;
    for (int f0 = 0; f0 < b->n_frames; f0 += G, ++gi) {
        const int nf = (b->n_frames - f0 < G) ? b->n_frames - f0 : G;
        const int l = gi % lanes;
        if ((s = filter_frames_graphed(ctx, b, f0, nf, g, p, roi_in, lanes > 1 ? ctx->lane_st[l] : st, (float*)((uint8_t*)ctx->ws + l * lane_ws)))) return s;
    }
    if (lanes > 1)
        for (int l = 0; l < lanes; ++l) {
            CK(cudaEventRecord(ctx->lane_ev[l], ctx->lane_st[l]));
            CK(cudaStreamWaitEvent(st, ctx->lane_ev[l], 0));
        }
    CK(call_leave(ctx, st));
    return DWLS_OK;
}

// ---------------------------------------------------------------------------------------------
// host-buffer entry points (what a cv::Mat facade calls): stage H2D, run, stage D2H, synchronise
// ---------------------------------------------------------------------------------------------
static bool img_ok(const dwls_image* m, int depth, int cn_a, int cn_b)
{
    return m && m->data && m->rows > 0 && m->cols > 0 && m->depth == depth && (m->channels == cn_a || m->channels == cn_b);
}
static size_t elem_size(int depth) { return depth == DWLS_8U ? 1 : depth == DWLS_16S ? 2 : 4; }

// ---------------------------------------------------------------------------------------------
// host <-> device transfers of the single-call entry points.  A job moves `rows` rows of `row_bytes` between caller memory
// (any row step, normally pageable) and a dense device array.  Small transfers go through the driver's pageable path on the
// launching stream; larger ones through pinned staging, split into row slabs over XFER_T host threads, each with its own
// copy stream, a slab moving in chunks so that the host-side memcpy of one chunk overlaps the DMA of another.
// ---------------------------------------------------------------------------------------------
struct XferJob { uint8_t* host; int64_t hstep; size_t row_bytes; int rows; uint8_t* dev; size_t pin_off; };
#ifndef DWLS_XFER_T
#define DWLS_XFER_T 4
#endif
static const int XFER_T = DWLS_XFER_T;          // host threads (and copy streams) of a staged transfer, at most 8
static const size_t XFER_MIN = (size_t)2 << 20, XFER_CHUNK = (size_t)1 << 20;

static dwls_status xfer_prepare(dwls_ctx* ctx, XferJob* jobs, int n, size_t* total)
{
    size_t off = 0;
    for (int i = 0; i < n; ++i) { jobs[i].pin_off = off; off += align256(jobs[i].row_bytes * jobs[i].rows); }
    *total = off;
    if (off < XFER_MIN) return DWLS_OK;
    if (ctx->hpin_bytes < off) {
        if (ctx->hpin) { cudaFreeHost(ctx->hpin); ctx->hpin = nullptr; ctx->hpin_bytes = 0; }
        CK(cudaHostAlloc(&ctx->hpin, off, cudaHostAllocDefault));
        ctx->hpin_bytes = off;
    }
    for (int t = 0; t < XFER_T; ++t) {
        if (!ctx->xs[t]) CK(cudaStreamCreateWithFlags(&ctx->xs[t], cudaStreamNonBlocking));
        for (int k = 0; k < 2; ++k) if (!ctx->xe[t][k]) CK(cudaEventCreateWithFlags(&ctx->xe[t][k], cudaEventDisableTiming));
    }
    if (!ctx->xready) CK(cudaEventCreateWithFlags(&ctx->xready, cudaEventDisableTiming));
    return DWLS_OK;
}

// rows [r0, r1) of a job between caller memory and the pinned buffer
static void xfer_rows(const XferJob& j, uint8_t* pin, int r0, int r1, bool to_pin)
{
    uint8_t* p = pin + j.pin_off + (size_t)r0 * j.row_bytes;
    uint8_t* h = j.host + (size_t)r0 * j.hstep;
    if (j.hstep == (int64_t)j.row_bytes) { if (to_pin) memcpy(p, h, (size_t)(r1 - r0) * j.row_bytes); else memcpy(h, p, (size_t)(r1 - r0) * j.row_bytes); return; }
    for (int r = r0; r < r1; ++r, p += j.row_bytes, h += j.hstep) { if (to_pin) memcpy(p, h, j.row_bytes); else memcpy(h, p, j.row_bytes); }
}

// host -> device; on return the copies are queued and `st` waits for them
static dwls_status xfer_upload(dwls_ctx* ctx, XferJob* jobs, int n, cudaStream_t st)
{
    size_t total = 0;
    dwls_status s = xfer_prepare(ctx, jobs, n, &total);
    if (s) return s;
    if (total < XFER_MIN) {
        for (int i = 0; i < n; ++i)
            CK(cudaMemcpy2DAsync(jobs[i].dev, jobs[i].row_bytes, jobs[i].host, (size_t)jobs[i].hstep, jobs[i].row_bytes, jobs[i].rows, cudaMemcpyHostToDevice, st));
        return DWLS_OK;
    }
    std::atomic<int> err(0);
    const std::function<void(int)> work = [&](int t) {
        if (cudaSetDevice(ctx->device) != cudaSuccess) { err = 1; return; }
        for (int i = 0; i < n; ++i) {
            const XferJob& j = jobs[i];
            const int r0 = (int)((int64_t)j.rows * t / XFER_T), r1 = (int)((int64_t)j.rows * (t + 1) / XFER_T);
            const int step = (int)std::max<size_t>(1, XFER_CHUNK / j.row_bytes);
            for (int r = r0; r < r1; r += step) {
                const int re = std::min(r + step, r1);
                xfer_rows(j, (uint8_t*)ctx->hpin, r, re, true);
                if (cudaMemcpyAsync(j.dev + (size_t)r * j.row_bytes, (uint8_t*)ctx->hpin + j.pin_off + (size_t)r * j.row_bytes, (size_t)(re - r) * j.row_bytes,
                                    cudaMemcpyHostToDevice, ctx->xs[t]) != cudaSuccess) { err = 1; return; }
            }
        }
        if (cudaEventRecord(ctx->xe[t][0], ctx->xs[t]) != cudaSuccess) err = 1;
    };
    ctx->pool.run(XFER_T, work);
    if (err) { cudaGetLastError(); return fail(ctx, DWLS_CUDA_ERROR, "host-to-device staging copy failed"); }
    for (int t = 0; t < XFER_T; ++t) CK(cudaStreamWaitEvent(st, ctx->xe[t][0], 0));
    return DWLS_OK;
}

// device -> host after everything queued on `st`; returns when the caller's buffers are filled
static dwls_status xfer_download(dwls_ctx* ctx, XferJob* jobs, int n, cudaStream_t st)
{
    size_t total = 0;
    dwls_status s = xfer_prepare(ctx, jobs, n, &total);
    if (s) return s;
    if (total < XFER_MIN) {
        for (int i = 0; i < n; ++i)
            CK(cudaMemcpy2DAsync(jobs[i].host, (size_t)jobs[i].hstep, jobs[i].dev, jobs[i].row_bytes, jobs[i].row_bytes, jobs[i].rows, cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        return DWLS_OK;
    }
    CK(cudaEventRecord(ctx->xready, st));
    std::atomic<int> err(0);
    const std::function<void(int)> work = [&](int t) {
        if (cudaSetDevice(ctx->device) != cudaSuccess || cudaStreamWaitEvent(ctx->xs[t], ctx->xready, 0) != cudaSuccess) { err = 1; return; }
        // the chunks of this thread's slabs, two in flight: chunk c+1 is being copied while chunk c is moved to the caller
        struct Chunk { int job, r0, r1; };
        std::vector<Chunk> ch;
        for (int i = 0; i < n; ++i) {
            const XferJob& j = jobs[i];
            const int r0 = (int)((int64_t)j.rows * t / XFER_T), r1 = (int)((int64_t)j.rows * (t + 1) / XFER_T);
            const int step = (int)std::max<size_t>(1, XFER_CHUNK / j.row_bytes);
            for (int r = r0; r < r1; r += step) ch.push_back({ i, r, std::min(r + step, r1) });
        }
        auto issue = [&](size_t c) {
            const XferJob& j = jobs[ch[c].job];
            const size_t off = (size_t)ch[c].r0 * j.row_bytes;
            if (cudaMemcpyAsync((uint8_t*)ctx->hpin + j.pin_off + off, j.dev + off, (size_t)(ch[c].r1 - ch[c].r0) * j.row_bytes, cudaMemcpyDeviceToHost, ctx->xs[t]) != cudaSuccess ||
                cudaEventRecord(ctx->xe[t][c & 1], ctx->xs[t]) != cudaSuccess) err = 1;
        };
        if (!ch.empty()) issue(0);
        for (size_t c = 0; c < ch.size() && !err; ++c) {
            if (cudaEventSynchronize(ctx->xe[t][c & 1]) != cudaSuccess) { err = 1; break; }
            if (c + 1 < ch.size()) issue(c + 1);
            xfer_rows(jobs[ch[c].job], (uint8_t*)ctx->hpin, ch[c].r0, ch[c].r1, false);
        }
    };
    ctx->pool.run(XFER_T, work);
    if (err) { cudaGetLastError(); return fail(ctx, DWLS_CUDA_ERROR, "device-to-host staging copy failed"); }
    CK(cudaStreamSynchronize(st));
    return DWLS_OK;
}

static dwls_status filter_host_impl(dwls_ctx* ctx, const dwls_image* dl, const dwls_image* dr, const dwls_image* view,
                                    const int* roi_in, const dwls_params* p, dwls_image* out, dwls_image* conf, int* roi_out,
                                    uint8_t* mask, int64_t mask_step, int mask_view_cols)
{
    if (!ctx) return DWLS_BAD_ARG;
    dwls_status s = check_params(ctx, p);
    if (s) return s;
    const bool mask_only = mask != nullptr;
    if (!img_ok(dl, DWLS_16S, 1, 1)) return fail(ctx, DWLS_BAD_ARG, "left disparity map must be a non-empty CV_16SC1 image");          // 221
    if (!mask_only && !img_ok(view, DWLS_8U, 1, 3)) return fail(ctx, DWLS_BAD_ARG, "left view must be a non-empty CV_8UC1 or CV_8UC3 image"); // 222
    const bool useconf = p->use_confidence != 0;
    if (useconf) {
        if (!img_ok(dr, DWLS_16S, 1, 1)) return fail(ctx, DWLS_BAD_ARG, "right disparity map must be a non-empty CV_16SC1 image");      // 262
        if (dr->cols != dl->cols || dr->rows != dl->rows) return fail(ctx, DWLS_BAD_ARG, "left and right disparity maps differ in size"); // 263-264
    }
    const int dw = dl->cols, dh = dl->rows;
    const int W = mask_only ? (mask_view_cols > 0 ? mask_view_cols : dw) : view->cols;
    const int H = mask_only ? (W == dw ? dh : (int)((int64_t)dh * W / dw)) : view->rows;
    const int vcn = mask_only ? 1 : view->channels;
    if (!mask_only) {
        if (!out || !out->data || out->rows != H || out->cols != W || out->depth != DWLS_16S || out->channels != 1)
            return fail(ctx, DWLS_BAD_ARG, "output must be a caller-allocated CV_16SC1 image of the view size");
        if (conf && conf->data && (conf->rows != H || conf->cols != W || conf->depth != DWLS_32F || conf->channels != 1))
            return fail(ctx, DWLS_BAD_ARG, "confidence output must be CV_32FC1 of the view size");
    }
    Geom g;
    s = make_geom(ctx, dw, dh, W, H, vcn, roi_in, p, &g);
    if (roi_out) { roi_out[0] = g.rx; roi_out[1] = g.ry; roi_out[2] = g.rw; roi_out[3] = g.rh; }
    if (s) return s;
    CK(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    CK(call_enter(ctx, st));
    const bool want_conf = !mask_only && useconf && conf && conf->data;
    const size_t b_dl = align256((size_t)dw * dh * 2), b_view = mask_only ? 0 : align256((size_t)W * H * vcn),
                 b_out = mask_only ? 0 : align256((size_t)W * H * 2), b_conf = want_conf ? align256((size_t)W * H * 4) : 0,
                 b_mask = mask_only ? align256((size_t)dw * dh) : 0;
    if ((s = ensure(ctx, &ctx->stage, &ctx->stage_bytes, 2 * b_dl + b_view + b_out + b_conf + b_mask))) return s;
    uint8_t* d = (uint8_t*)ctx->stage;
    uint8_t* d_dl = d; uint8_t* d_dr = d_dl + b_dl; uint8_t* d_view = d_dr + b_dl; uint8_t* d_out = d_view + b_view;
    uint8_t* d_conf = d_out + b_out; uint8_t* d_mask = d_conf + b_conf;
    ctx->h2d_bytes += (int64_t)dw * dh * 2 * (useconf ? 2 : 1) + (mask_only ? 0 : (int64_t)W * H * vcn);
    ctx->d2h_bytes += mask_only ? (int64_t)dw * dh : (int64_t)W * H * 2 + (want_conf ? (int64_t)W * H * 4 : 0);
    XferJob in[3], outj[2];
    int n_in = 0, n_out = 0;
    in[n_in++] = { (uint8_t*)dl->data, dl->step, (size_t)dw * 2, dh, d_dl, 0 };
    if (useconf) in[n_in++] = { (uint8_t*)dr->data, dr->step, (size_t)dw * 2, dh, d_dr, 0 };
    if (!mask_only) {
        in[n_in++] = { (uint8_t*)view->data, view->step, (size_t)W * vcn, H, d_view, 0 };
        outj[n_out++] = { (uint8_t*)out->data, out->step, (size_t)W * 2, H, d_out, 0 };
        if (want_conf) outj[n_out++] = { (uint8_t*)conf->data, conf->step, (size_t)W * 4, H, d_conf, 0 };
        size_t reserve = 0;
        if ((s = xfer_prepare(ctx, outj, n_out, &reserve))) return s;      // the pinned buffer must not grow between upload and download
    }
    if ((s = xfer_upload(ctx, in, n_in, st))) return s;
    dwls_batch b;
    memset(&b, 0, sizeof b);
    b.n_frames = 1;
    b.disp_left = (const int16_t*)d_dl; b.dl_step = (int64_t)dw * 2;
    b.disp_right = useconf ? (const int16_t*)d_dr : nullptr; b.dr_step = (int64_t)dw * 2;
    b.disp_cols = dw; b.disp_rows = dh;
    b.view = d_view; b.view_step = (int64_t)W * vcn; b.view_cols = W; b.view_rows = H; b.view_channels = vcn;
    b.out = (int16_t*)d_out; b.out_step = (int64_t)W * 2;
    b.conf = want_conf ? (float*)d_conf : nullptr; b.conf_step = (int64_t)W * 4;
    const WsLayout L = ws_layout(g, useconf);
    if ((s = ensure(ctx, &ctx->ws, &ctx->ws_bytes, align256(L.per_frame * sizeof(float))))) return s;
    if (mask_only) {
        if (!useconf) return fail(ctx, DWLS_BAD_ARG, "the LRC mask exists only when filtering with confidence");
        if ((s = filter_frames(ctx, &b, 0, 1, g, p, d_mask, dw, true, st))) return s;
        CK(cudaMemcpy2DAsync(mask, mask_step, d_mask, dw, dw, dh, cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        return DWLS_OK;
    }
    if ((s = ensure_lut(ctx, (float)p->sigma_color, st))) return s;
    if ((s = filter_frames_graphed(ctx, &b, 0, 1, g, p, roi_in, st))) return s;
    return xfer_download(ctx, outj, n_out, st);
}

extern "C" dwls_status dwls_filter_host(dwls_ctx* ctx, const dwls_image* dl, const dwls_image* dr, const dwls_image* view,
                                        const int* roi_in, const dwls_params* p, dwls_image* out, dwls_image* conf, int* roi_out)
{
    return filter_host_impl(ctx, dl, dr, view, roi_in, p, out, conf, roi_out, nullptr, 0, 0);
}

extern "C" dwls_status dwls_lrc_mask_host(dwls_ctx* ctx, const dwls_image* dl, const dwls_image* dr, int view_cols, const int* roi_in,
                                          const dwls_params* p, uint8_t* mask, int64_t mask_step)
{
    if (!mask) return fail(ctx, DWLS_BAD_ARG, "mask is NULL");
    return filter_host_impl(ctx, dl, dr, nullptr, roi_in, p, nullptr, nullptr, nullptr, mask, mask_step, view_cols);
}

// ---------------------------------------------------------------------------------------------
// fastGlobalSmootherFilter (fgs_filter.cpp:687-691)
// ---------------------------------------------------------------------------------------------
static dwls_status fgs_check(dwls_ctx* ctx, const dwls_image* guide, const dwls_image* src, const dwls_image* dst,
                             double lambda, double sigma, int num_iter)
{
    if (!guide || !guide->data || guide->rows <= 0 || guide->cols <= 0 || !(lambda >= 0) || !(sigma >= 0) || num_iter < 1)
        return fail(ctx, DWLS_BAD_ARG, "guide must be non-empty, lambda >= 0, sigma_color >= 0, num_iter >= 1");                       // 143
    if (guide->depth != DWLS_8U || (guide->channels != 1 && guide->channels != 3))
        return fail(ctx, DWLS_BAD_ARG, "guide must be CV_8UC1 or CV_8UC3");                                                              // 144
    if (!src || !src->data || src->rows <= 0 || src->cols <= 0 || (src->depth != DWLS_8U && src->depth != DWLS_16S && src->depth != DWLS_32F) ||
        src->channels < 1 || src->channels > 4)
        return fail(ctx, DWLS_BAD_ARG, "src must be non-empty CV_8U / CV_16S / CV_32F with at most 4 channels");                        // 184
    if (src->rows != guide->rows || src->cols != guide->cols)
        return fail(ctx, DWLS_BAD_SIZE, "Size of the filtered image must be equal to the size of the guide image");                     // 185-189
    if (!dst || !dst->data || dst->rows != src->rows || dst->cols != src->cols || dst->depth != src->depth || dst->channels != src->channels)
        return fail(ctx, DWLS_BAD_ARG, "dst must be caller-allocated with the size and type of src");
    return DWLS_OK;
}

template <typename T>
static void launch_unpack(const dwls_image* src, float* planes, size_t plane, int pitch, cudaStream_t st)
{
    const dim3 grid((src->cols + 1023) / 1024, src->rows);
    const T* p = (const T*)src->data;
    switch (src->channels) {
    case 1: unpack_kernel<T, 1><<<grid, 256, 0, st>>>(p, src->step, src->cols, planes, plane, pitch); break;
    case 2: unpack_kernel<T, 2><<<grid, 256, 0, st>>>(p, src->step, src->cols, planes, plane, pitch); break;
    case 3: unpack_kernel<T, 3><<<grid, 256, 0, st>>>(p, src->step, src->cols, planes, plane, pitch); break;
    default: unpack_kernel<T, 4><<<grid, 256, 0, st>>>(p, src->step, src->cols, planes, plane, pitch); break;
    }
}
template <typename T>
static void launch_pack(const float* planes, size_t plane, int pitch, dwls_image* dst, cudaStream_t st)
{
    const dim3 grid((dst->cols + 1023) / 1024, dst->rows);
    T* p = (T*)dst->data;
    switch (dst->channels) {
    case 1: pack_kernel<T, 1><<<grid, 256, 0, st>>>(planes, plane, pitch, dst->cols, p, dst->step); break;
    case 2: pack_kernel<T, 2><<<grid, 256, 0, st>>>(planes, plane, pitch, dst->cols, p, dst->step); break;
    case 3: pack_kernel<T, 3><<<grid, 256, 0, st>>>(planes, plane, pitch, dst->cols, p, dst->step); break;
    default: pack_kernel<T, 4><<<grid, 256, 0, st>>>(planes, plane, pitch, dst->cols, p, dst->step); break;
    }
}

// workspace: chor | cvert | up to 4 planes (+ optional export of the weights)
static dwls_status fgs_dev_impl(dwls_ctx* ctx, const dwls_image* guide, const dwls_image* src, dwls_image* dst, double lambda,
                                double sigma, double att, int num_iter, cudaStream_t st, float* chor_out, float* cvert_out)
{
    const int w = guide->cols, h = guide->rows, pitch = round_up(w, 8);
    SolvePlan pl;
    if (!plan_solver(w, h, &pl)) return fail(ctx, DWLS_UNSUPPORTED, "images larger than 8192 x 8192 are not supported by the on-chip solver");
    const size_t plane = (size_t)pitch * h;
    const int S = src ? src->channels : 0;
    dwls_status s;
    TPass T;
    T.tplane = pl.vT ? (size_t)pl.tpitch * w : 0;
    if ((s = ensure(ctx, &ctx->ws, &ctx->ws_bytes, align256(((2 + S) * plane + 3 * T.tplane) * sizeof(float))))) return s;
    if ((s = ensure_lut(ctx, (float)sigma, st))) return s;
    float* chor = (float*)ctx->ws; float* cvert = chor + plane; float* planes = cvert + plane;
    T.cvertT = planes + (size_t)S * plane; T.pT[0] = T.cvertT + T.tplane; T.pT[1] = T.pT[0] + T.tplane;
    if ((s = launch_weights(ctx, (const uint8_t*)guide->data, guide->step, 0, guide->channels, w, h, pitch, chor, cvert, plane, 1, st, chor_out ? 0 : pl.wstrip))) return s;   // the exported planes are row-major
    if (chor_out) {
        CK(cudaMemcpy2DAsync(chor_out, (size_t)w * 4, chor, (size_t)pitch * 4, (size_t)w * 4, h, cudaMemcpyDeviceToDevice, st));
        CK(cudaMemcpy2DAsync(cvert_out, (size_t)w * 4, cvert, (size_t)pitch * 4, (size_t)w * 4, h, cudaMemcpyDeviceToDevice, st));
        return DWLS_OK;
    }
    // CV_32FC1 with dense, 32-byte aligned rows of a multiple of 8 pixels IS a solver plane: the first row pass reads src and
    // writes dst, every later pass works in place on dst, and no conversion kernels run
    const bool direct = S == 1 && src->depth == DWLS_32F && pitch == w && src->step == (int64_t)w * 4 && dst->step == (int64_t)w * 4 &&
                        (((uintptr_t)src->data | (uintptr_t)dst->data) & 31) == 0;
    if (!direct) {
        KTimer kt_(ctx, K_UNPACK, st);
        if (src->depth == DWLS_8U) launch_unpack<uint8_t>(src, planes, plane, pitch, st);
        else if (src->depth == DWLS_16S) launch_unpack<int16_t>(src, planes, plane, pitch, st);
        else launch_unpack<float>(src, planes, plane, pitch, st);
    }
    CK(cudaGetLastError());
    if (pl.vT && (s = transpose_cvert(ctx, cvert, plane, pitch, w, h, pl, T, 1, st))) return s;
    // channels are independent solves sharing the weights (fgs_filter.cpp:200-221): two planes per launch
    for (int c0 = 0; c0 < S; c0 += 2) {
        const int ns = (S - c0 >= 2) ? 2 : 1;
        SolveArgs a;
        memset(&a, 0, sizeof a);
        a.w_frame = plane; a.p_frame = plane; a.pitch = pitch; a.w = w; a.h = h;
        for (int k = 0; k < ns; ++k) { a.in[k] = planes + (size_t)(c0 + k) * plane; a.out[k] = planes + (size_t)(c0 + k) * plane; }
        if (direct) { a.in[0] = (const float*)src->data; a.out[0] = (float*)dst->data; }
        if ((s = run_sweeps(ctx, a, chor, cvert, ns, FIN_FLOAT, pl, 1, lambda, att, num_iter, st, &T))) return s;
    }
    if (direct) return DWLS_OK;
    { KTimer kt_(ctx, K_PACK, st);
    if (dst->depth == DWLS_8U) launch_pack<uint8_t>(planes, plane, pitch, dst, st);
    else if (dst->depth == DWLS_16S) launch_pack<int16_t>(planes, plane, pitch, dst, st);
    else launch_pack<float>(planes, plane, pitch, dst, st); }
    CK(cudaGetLastError());
    return DWLS_OK;
}

extern "C" dwls_status dwls_fgs_dev(dwls_ctx* ctx, const dwls_image* guide, const dwls_image* src, dwls_image* dst, double lambda,
                                    double sigma, double att, int num_iter, void* stream)
{
    if (!ctx) return DWLS_BAD_ARG;
    dwls_status s = fgs_check(ctx, guide, src, dst, lambda, sigma, num_iter);
    if (s) return s;
    CK(cudaSetDevice(ctx->device));
    cudaStream_t st = stream ? (cudaStream_t)stream : ctx->stream;
    CK(call_enter(ctx, st));
    if ((s = fgs_dev_impl(ctx, guide, src, dst, lambda, sigma, att, num_iter, st, nullptr, nullptr))) return s;
    CK(call_leave(ctx, st));
    return DWLS_OK;
}

extern "C" dwls_status dwls_fgs_host(dwls_ctx* ctx, const dwls_image* guide, const dwls_image* src, dwls_image* dst, double lambda,
                                     double sigma, double att, int num_iter)
{
    if (!ctx) return DWLS_BAD_ARG;
    dwls_status s = fgs_check(ctx, guide, src, dst, lambda, sigma, num_iter);
    if (s) return s;
    CK(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    CK(call_enter(ctx, st));
    const int w = guide->cols, h = guide->rows;
    const size_t grow = (size_t)w * guide->channels, srow = (size_t)w * src->channels * elem_size(src->depth);
    const size_t b_g = align256(grow * h), b_s = align256(srow * h);
    if ((s = ensure(ctx, &ctx->stage, &ctx->stage_bytes, b_g + 2 * b_s))) return s;
    uint8_t* d_g = (uint8_t*)ctx->stage; uint8_t* d_s = d_g + b_g; uint8_t* d_d = d_s + b_s;
    ctx->h2d_bytes += (int64_t)(grow + srow) * h;
    ctx->d2h_bytes += (int64_t)srow * h;
    XferJob in[2] = { { (uint8_t*)guide->data, guide->step, grow, h, d_g, 0 }, { (uint8_t*)src->data, src->step, srow, h, d_s, 0 } };
    XferJob outj[1] = { { (uint8_t*)dst->data, dst->step, srow, h, d_d, 0 } };
    if ((s = xfer_upload(ctx, in, 2, st))) return s;          // in >= out: the pinned buffer does not grow again for the download
    dwls_image G = *guide, Sm = *src, D = *dst;
    G.data = d_g; G.step = (int64_t)grow; Sm.data = d_s; Sm.step = (int64_t)srow; D.data = d_d; D.step = (int64_t)srow;
    if ((s = fgs_dev_impl(ctx, &G, &Sm, &D, lambda, sigma, att, num_iter, st, nullptr, nullptr))) return s;
    return xfer_download(ctx, outj, 1, st);
}

extern "C" dwls_status dwls_fgs_weights_host(dwls_ctx* ctx, const dwls_image* guide, double sigma, float* chor, float* cvert)
{
    if (!ctx) return DWLS_BAD_ARG;
    if (!guide || !guide->data || guide->rows <= 0 || guide->cols <= 0 || guide->depth != DWLS_8U || (guide->channels != 1 && guide->channels != 3) ||
        !(sigma >= 0) || !chor || !cvert)
        return fail(ctx, DWLS_BAD_ARG, "guide must be non-empty CV_8UC1 / CV_8UC3, sigma_color >= 0");
    CK(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    CK(call_enter(ctx, st));
    const int w = guide->cols, h = guide->rows;
    const size_t grow = (size_t)w * guide->channels, b_g = align256(grow * h), b_w = align256((size_t)w * h * 4);
    dwls_status s;
    if ((s = ensure(ctx, &ctx->stage, &ctx->stage_bytes, b_g + 2 * b_w))) return s;
    uint8_t* d_g = (uint8_t*)ctx->stage; float* d_h = (float*)(d_g + b_g); float* d_v = (float*)(d_g + b_g + b_w);
    CK(cudaMemcpy2DAsync(d_g, grow, guide->data, guide->step, grow, h, cudaMemcpyHostToDevice, st));
    dwls_image G = *guide;
    G.data = d_g; G.step = (int64_t)grow;
    if ((s = fgs_dev_impl(ctx, &G, nullptr, nullptr, 0.0, sigma, 0.25, 1, st, d_h, d_v))) return s;
    CK(cudaMemcpyAsync(chor, d_h, (size_t)w * h * 4, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(cvert, d_v, (size_t)w * h * 4, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    return DWLS_OK;
}

// nf frames of `rows` rows of `row_bytes` between host and device: one transfer when both sides are dense or when the rows
// of all frames are equally spaced on both sides (a column window of a dense batch), else one per frame.
// dst_step < 0: the destination rows are dense (device staging).
static dwls_status copy_frames(dwls_ctx* ctx, void* dst, int64_t dst_frame, const void* src, int64_t src_frame, int64_t src_step,
                               size_t row_bytes, int rows, int nf, cudaMemcpyKind kind, cudaStream_t st, int64_t dst_step = -1)
{
    if (dst_step < 0) dst_step = (int64_t)row_bytes;
    (kind == cudaMemcpyHostToDevice ? ctx->h2d_bytes : ctx->d2h_bytes) += (int64_t)row_bytes * rows * nf;
    const bool dense_rows = src_step == (int64_t)row_bytes && dst_step == (int64_t)row_bytes;
    const int64_t frame_bytes = (int64_t)row_bytes * rows;
    if (dense_rows && (nf == 1 || (src_frame == frame_bytes && dst_frame == frame_bytes))) {
        CK(cudaMemcpyAsync(dst, src, (size_t)frame_bytes * nf, kind, st));
        return DWLS_OK;
    }
    if (nf == 1 || (src_frame == src_step * rows && dst_frame == dst_step * rows)) {
        CK(cudaMemcpy2DAsync(dst, (size_t)dst_step, src, (size_t)src_step, row_bytes, (size_t)rows * nf, kind, st));
        return DWLS_OK;
    }
    for (int f = 0; f < nf; ++f) {
        uint8_t* d = (uint8_t*)dst + (size_t)f * dst_frame;
        const uint8_t* s = (const uint8_t*)src + (size_t)f * src_frame;
        if (dense_rows) CK(cudaMemcpyAsync(d, s, (size_t)frame_bytes, kind, st));
        else CK(cudaMemcpy2DAsync(d, (size_t)dst_step, s, (size_t)src_step, row_bytes, rows, kind, st));
    }
    return DWLS_OK;
}

// the part of an input the kernels read: [lo - margin, lo + len + margin) clipped to [0, full), widened to multiples of
// `align` elements.  The margin covers the bilinear taps of the half-resolution mode (disparity_filters.cpp:268-277);
// staged cells outside the window keep whatever an earlier call left there and are never used.
static void read_window(int lo, int len, int full, int margin, int align, int* w0, int* w1)
{
    int a = lo - margin, b = lo + len + margin;
    if (a < 0) a = 0;
    if (b > full) b = full;
    a = a / align * align;
    b = (b + align - 1) / align * align;
    if (b > full) b = full;
    *w0 = a; *w1 = b;
}

// ---------------------------------------------------------------------------------------------
// batch of frames in HOST memory (pinned for full speed): two lanes, each with its own stream,
// staging area and workspace, so the H2D copy of one group overlaps the kernels and the D2H copy
// of the previous one.  Returns after everything has landed in the host outputs.
// ---------------------------------------------------------------------------------------------
extern "C" dwls_status dwls_filter_batch_host(dwls_ctx* ctx, const dwls_batch* hb, const int* roi_in, const dwls_params* p, int* roi_out)
{
    if (!ctx) return DWLS_BAD_ARG;
    dwls_status s = check_params(ctx, p);
    if (s) return s;
    if ((s = validate_batch(ctx, hb, p))) return s;
    Geom g;
    s = make_geom(ctx, hb->disp_cols, hb->disp_rows, hb->view_cols, hb->view_rows, hb->view_channels, roi_in, p, &g);
    if (roi_out) { roi_out[0] = g.rx; roi_out[1] = g.ry; roi_out[2] = g.rw; roi_out[3] = g.rh; }
    if (s) return s;
    CK(cudaSetDevice(ctx->device));
    if (!ctx->stream2) CK(cudaStreamCreateWithFlags(&ctx->stream2, cudaStreamNonBlocking));
    CK(call_enter(ctx, ctx->stream));
    CK(call_enter(ctx, ctx->stream2));
    const bool useconf = p->use_confidence != 0, want_conf = useconf && hb->conf;
    const int dw = g.dw, dh = g.dh, W = g.W, H = g.H, vcn = g.vcn;
    const WsLayout L = ws_layout(g, useconf);
    int G = pick_group(ctx, L.per_frame * sizeof(float), hb->n_frames);
    if (ctx->group_frames == 0 && G > 8) G = 8;           // smaller groups pipeline better against PCIe (the path is H2D-bound)
    const size_t f_dl = align256((size_t)dw * dh * 2), f_view = align256((size_t)W * H * vcn), f_out = align256((size_t)W * H * 2),
                 f_conf = want_conf ? align256((size_t)W * H * 4) : 0;
    const size_t lane_stage = (size_t)G * (2 * f_dl + f_view + f_out + f_conf);
    const size_t lane_ws = align256((size_t)G * L.per_frame * sizeof(float));
    if ((s = ensure(ctx, &ctx->stage, &ctx->stage_bytes, 2 * lane_stage))) return s;
    if ((s = ensure(ctx, &ctx->ws, &ctx->ws_bytes, 2 * lane_ws))) return s;
    if ((s = ensure_lut(ctx, (float)p->sigma_color, ctx->stream))) return s;
    CK(cudaStreamSynchronize(ctx->stream));
    cudaStream_t lanes[2] = { ctx->stream, ctx->stream2 };
    int lane = 0;
    // the H2D engine is the bottleneck and stays busy back to back; what is exposed is the tail (kernels + D2H of the
    // last group), so the last G frames go out in halving groups (.., G/2, .., 1, 1)
    for (int f0 = 0, nf = 0; f0 < hb->n_frames; f0 += nf, lane ^= 1) {
        const int left = hb->n_frames - f0;
        nf = left > G ? G : (left + 1) / 2;
        cudaStream_t st = lanes[lane];
        uint8_t* d = (uint8_t*)ctx->stage + lane * lane_stage;
        uint8_t* d_dl = d; uint8_t* d_dr = d_dl + (size_t)G * f_dl; uint8_t* d_view = d_dr + (size_t)G * f_dl;
        uint8_t* d_out = d_view + (size_t)G * f_view; uint8_t* d_conf = d_out + (size_t)G * f_out;
        {
            // upload only what the filter reads (the windows are whole frames when the ROI is the whole image)
            dwls_status cs;
            int x0, x1, y0, y1;
            // DWLS_E2E_WINDOWS=0 (A/B runs): whole frames in one contiguous copy each instead of the ROI column windows
            static const bool windows = [] { const char* e = getenv("DWLS_E2E_WINDOWS"); return !(e && e[0] == '0'); }();
            const int wmargin = windows ? 8 : (1 << 20);
            read_window(g.ry, g.rh, dh, 8, 1, &y0, &y1);
            if (y1 - y0 != dh) { y0 = 0; y1 = dh; }           // row windows would split the batch into one copy per frame: not worth it
            read_window(g.rx, g.rw, dw, wmargin, 32, &x0, &x1);
            if ((cs = copy_frames(ctx, d_dl + (size_t)x0 * 2, f_dl, (const uint8_t*)hb->disp_left + (size_t)f0 * hb->dl_frame + (size_t)x0 * 2, hb->dl_frame,
                                  hb->dl_step, (size_t)(x1 - x0) * 2, dh, nf, cudaMemcpyHostToDevice, st, (int64_t)dw * 2))) return cs;
            read_window(useconf ? g.rrx : 0, g.rw, dw, wmargin, 32, &x0, &x1);
            if (useconf && (cs = copy_frames(ctx, d_dr + (size_t)x0 * 2, f_dl, (const uint8_t*)hb->disp_right + (size_t)f0 * hb->dr_frame + (size_t)x0 * 2,
                                             hb->dr_frame, hb->dr_step, (size_t)(x1 - x0) * 2, dh, nf, cudaMemcpyHostToDevice, st, (int64_t)dw * 2))) return cs;
            read_window(g.fx, g.fw, W, wmargin, 32, &x0, &x1);
            if ((cs = copy_frames(ctx, d_view + (size_t)x0 * vcn, f_view, hb->view + (size_t)f0 * hb->view_frame + (size_t)x0 * vcn, hb->view_frame, hb->view_step,
                                  (size_t)(x1 - x0) * vcn, H, nf, cudaMemcpyHostToDevice, st, (int64_t)W * vcn))) return cs;
        }
        dwls_batch b;
        memset(&b, 0, sizeof b);
        b.n_frames = nf;
        b.disp_left = (const int16_t*)d_dl; b.dl_step = (int64_t)dw * 2; b.dl_frame = (int64_t)f_dl;
        b.disp_right = useconf ? (const int16_t*)d_dr : nullptr; b.dr_step = (int64_t)dw * 2; b.dr_frame = (int64_t)f_dl;
        b.disp_cols = dw; b.disp_rows = dh;
        b.view = d_view; b.view_step = (int64_t)W * vcn; b.view_frame = (int64_t)f_view; b.view_cols = W; b.view_rows = H; b.view_channels = vcn;
        b.out = (int16_t*)d_out; b.out_step = (int64_t)W * 2; b.out_frame = (int64_t)f_out;
        b.conf = want_conf ? (float*)d_conf : nullptr; b.conf_step = (int64_t)W * 4; b.conf_frame = (int64_t)f_conf;
        if ((s = filter_frames(ctx, &b, 0, nf, g, p, nullptr, 0, false, st, (float*)((uint8_t*)ctx->ws + lane * lane_ws)))) return s;
        {
            dwls_status cs;
            if ((cs = copy_frames(ctx, (uint8_t*)hb->out + (size_t)f0 * hb->out_frame, hb->out_frame, d_out, f_out, (int64_t)W * 2, (size_t)W * 2, H, nf, cudaMemcpyDeviceToHost, st, hb->out_step))) return cs;
            if (want_conf && (cs = copy_frames(ctx, (uint8_t*)hb->conf + (size_t)f0 * hb->conf_frame, hb->conf_frame, d_conf, f_conf, (int64_t)W * 4, (size_t)W * 4, H, nf, cudaMemcpyDeviceToHost, st, hb->conf_step))) return cs;
        }
    }
    CK(cudaStreamSynchronize(ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream2));
    ctx->busy_valid = false;                 // nothing of this ctx is in flight any more
    return DWLS_OK;
}


// ---------------------------------------------------------------------------------------------
// stereo block matching on the device: the producer of the filter's inputs.
// Reference call sites: samples/disparity_filtering.cpp:148-160 (StereoBM::create(max_disp, wsize), createRightMatcher,
// compute(left, right) / compute(right, left)) under the parameter contract of src/disparity_filters.cpp:389-403, 419-428.
// ---------------------------------------------------------------------------------------------
extern "C" void dwls_bm_default_params(dwls_bm_params* p)
{
    if (!p) return;
    p->num_disparities = 64; p->block_size = 21; p->min_disparity = 0;       // cv::StereoBM::create(0, 21)
    p->pre_filter_cap = 31; p->texture_threshold = 10; p->uniqueness_ratio = 15;
}

static dwls_status bm_check(dwls_ctx* ctx, const dwls_bm_params* p, int w, int h, int cn)
{
    if (!p) return fail(ctx, DWLS_BAD_ARG, "matcher params is NULL");
    if (cn != 1 && cn != 3) return fail(ctx, DWLS_BAD_ARG, "views must be CV_8UC1 (or CV_8UC3, converted like cvtColor BGR2GRAY)");
    if (p->num_disparities <= 0 || p->num_disparities % 16) return fail(ctx, DWLS_BAD_ARG, "numDisparities must be positive and divisible by 16");   // StereoBM::compute
    if (p->block_size < 5 || p->block_size > 255 || !(p->block_size & 1) || p->block_size >= (w < h ? w : h))
        return fail(ctx, DWLS_BAD_ARG, "SADWindowSize must be odd, be within 5..255 and be not larger than image width or height");
    if (p->pre_filter_cap < 1 || p->pre_filter_cap > 63) return fail(ctx, DWLS_BAD_ARG, "preFilterCap must be within 1..63");
    if (p->min_disparity > 0) return fail(ctx, DWLS_UNSUPPORTED, "minDisparity > 0 is not supported by the device matcher (OpenCV itself writes past the row end there)");
    if (p->texture_threshold < 0 || p->uniqueness_ratio < 0) return fail(ctx, DWLS_BAD_ARG, "texture threshold and uniqueness ratio must be non-negative");
    if (p->block_size * p->block_size * 2 * p->pre_filter_cap >= 65536) return fail(ctx, DWLS_UNSUPPORTED, "window too large for the 16-bit SAD accumulators of the device matcher");
    return DWLS_OK;
}

// prefilter both views, then match: view 0 = matcher(left, right) with p->min_disparity; view 1 (when disp_right) = the
// matcher createRightMatcher builds: compute(right, left) with minDisparity -(min + num) + 1 (disparity_filters.cpp:424)
static dwls_status bm_frames(dwls_ctx* ctx, const uint8_t* left, int64_t l_step, int64_t l_frame, const uint8_t* right, int64_t r_step, int64_t r_frame,
                             int cn, int w, int h, int frames, const dwls_bm_params* p, int16_t* dl, int64_t dl_step, int64_t dl_frame,
                             int16_t* dr, int64_t dr_step, int64_t dr_frame, cudaStream_t st)
{
    dwls_status s;
    const size_t plane = align256((size_t)w * h);
    if ((s = ensure(ctx, &ctx->bmws, &ctx->bmws_bytes, 2 * plane * frames))) return s;
    uint8_t* pl = (uint8_t*)ctx->bmws; uint8_t* pr = pl + plane * frames;
    PrefArgs pa;
    memset(&pa, 0, sizeof pa);
    pa.src[0] = left; pa.s_step[0] = l_step; pa.s_frame[0] = l_frame; pa.src[1] = right; pa.s_step[1] = r_step; pa.s_frame[1] = r_frame; pa.cn = cn;
    pa.dst[0] = pl; pa.dst[1] = pr; pa.d_step = w; pa.d_frame = (int64_t)plane; pa.w = w; pa.h = h; pa.cap = p->pre_filter_cap;
    TIMED(K_BM_PREF, st, bm_prefilter_kernel<<<dim3((w + 255) / 256, h, 2 * frames), 256, 0, st>>>(pa));
    CK(cudaGetLastError());
    BmArgs a0, a1;
    memset(&a0, 0, sizeof a0);
    a0.left = pl; a0.l_step = w; a0.l_frame = (int64_t)plane; a0.right = pr; a0.r_step = w; a0.r_frame = (int64_t)plane;
    a0.disp = dl; a0.d_step = dl_step; a0.d_frame = dl_frame; a0.w = w; a0.h = h;
    a0.ndisp = p->num_disparities; a0.wsz = p->block_size; a0.mindisp = p->min_disparity; a0.ftzero = p->pre_filter_cap;
    a0.texture = p->texture_threshold; a0.uniq = p->uniqueness_ratio;
    a1 = a0;
    const int views = dr ? 2 : 1;
    if (dr) {
        a1.left = pr; a1.right = pl; a1.disp = dr; a1.d_step = dr_step; a1.d_frame = dr_frame;
        a1.mindisp = -(p->min_disparity + p->num_disparities) + 1;
    }
    const BmPlan P0 = bm_plan(w, h, a0.ndisp, a0.wsz, a0.mindisp), P1 = bm_plan(w, h, a1.ndisp, a1.wsz, a1.mindisp);
    // frames * 2 blocks in z even for one view (the odd ones return at once): keeps the view / frame decoding in one place
    TIMED(K_BM_FILL, st, bm_fill_kernel<<<dim3((w + 255) / 256, h, 2 * frames), 256, 0, st>>>(a0, dr ? a1 : a0));
    CK(cudaGetLastError());
    const bool run0 = P0.lofs < w && P0.rofs < w && P0.width1 >= 1 && P0.y1 > P0.y0, run1 = dr && P1.lofs < w && P1.rofs < w && P1.width1 >= 1 && P1.y1 > P1.y0;
    if (!run0 && !run1) return DWLS_OK;                                     // numDisparities wider than the image: everything is FILTERED
    if ((run0 && !P0.ok) || (run1 && !P1.ok)) return fail(ctx, DWLS_UNSUPPORTED, "numDisparities x window too large for the device matcher's per-thread sums");
    const int width1 = std::max(run0 ? P0.width1 : 0, run1 ? P1.width1 : 0);
    const size_t smem = std::max(P0.smem, P1.smem);
    if (smem > 200 * 1024) return fail(ctx, DWLS_UNSUPPORTED, "numDisparities x window too large for the device matcher's shared memory");
    // the window size is a template parameter of the kernel (unrolled taps); sizes without an instantiation take the looped one
    const int need_g = std::max(run0 ? (P0.wc * (a0.ndisp / 4) + BM_T - 1) / BM_T : 0, run1 ? (P1.wc * (a1.ndisp / 4) + BM_T - 1) / BM_T : 0);
    const bool small = need_g <= 8;
    void (*kern)(const BmArgs, const BmArgs) = small ? bm_kernel<0, 8> : bm_kernel<0, BM_MAXG>;
#define DWLS_BM_CASE(W) case W: kern = small ? bm_kernel<W, 8> : bm_kernel<W, BM_MAXG>; break;
    switch (p->block_size) {
    DWLS_BM_CASE(5) DWLS_BM_CASE(7) DWLS_BM_CASE(9) DWLS_BM_CASE(11) DWLS_BM_CASE(13) DWLS_BM_CASE(15) DWLS_BM_CASE(17) DWLS_BM_CASE(19) DWLS_BM_CASE(21)
    default: break;
    }
#undef DWLS_BM_CASE
    if (ctx->smem_attr_done.insert((const void*)kern).second || smem > 48 * 1024)
        CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)std::max(smem, (size_t)48 * 1024)));
    if (!run0) { a0 = a1; }                                                  // (cannot happen for a left matcher with a valid right one; kept total)
    (void)views;
    TIMED(K_BM, st, kern<<<dim3((width1 + BM_TW - 1) / BM_TW, (P0.y1 - P0.y0 + BM_RB - 1) / BM_RB, 2 * frames), BM_T, smem, st>>>(a0, run1 ? a1 : a0));
    CK(cudaGetLastError());
    return DWLS_OK;
}

extern "C" dwls_status dwls_stereo_bm_dev(dwls_ctx* ctx, const dwls_views* v, const dwls_bm_params* p, int16_t* disp_left, int64_t dl_step, int64_t dl_frame,
                                          int16_t* disp_right, int64_t dr_step, int64_t dr_frame, void* stream)
{
    if (!ctx) return DWLS_BAD_ARG;
    if (!v || v->n_frames <= 0 || !v->left || !v->right || v->cols <= 0 || v->rows <= 0 || !disp_left) return fail(ctx, DWLS_BAD_ARG, "views / output missing");
    dwls_status s = bm_check(ctx, p, v->cols, v->rows, v->channels);
    if (s) return s;
    if ((dl_step | dl_frame | dr_step | dr_frame) & 1) return fail(ctx, DWLS_BAD_ARG, "CV_16S strides must be even");
    CK(cudaSetDevice(ctx->device));
    cudaStream_t st = stream ? (cudaStream_t)stream : ctx->stream;
    CK(call_enter(ctx, st));
    if ((s = bm_frames(ctx, v->left, v->left_step, v->left_frame, v->right, v->right_step, v->right_frame, v->channels, v->cols, v->rows, v->n_frames, p,
                       disp_left, dl_step / 2, dl_frame / 2, disp_right, dr_step / 2, dr_frame / 2, st))) return s;
    CK(call_leave(ctx, st));
    return DWLS_OK;
}

extern "C" dwls_status dwls_stereo_bm_host(dwls_ctx* ctx, const dwls_image* left, const dwls_image* right, const dwls_bm_params* p,
                                           dwls_image* disp_left, dwls_image* disp_right)
{
    if (!ctx) return DWLS_BAD_ARG;
    if (!img_ok(left, DWLS_8U, 1, 3) || !img_ok(right, DWLS_8U, 1, 3) || left->rows != right->rows || left->cols != right->cols || left->channels != right->channels)
        return fail(ctx, DWLS_BAD_ARG, "left and right views must be non-empty CV_8UC1 / CV_8UC3 images of equal size and type");    // StereoBM::compute
    const int w = left->cols, h = left->rows, cn = left->channels;
    dwls_status s = bm_check(ctx, p, w, h, cn);
    if (s) return s;
    if (!img_ok(disp_left, DWLS_16S, 1, 1) || disp_left->rows != h || disp_left->cols != w) return fail(ctx, DWLS_BAD_ARG, "disp_left must be a caller-allocated CV_16SC1 image of the view size");
    if (disp_right && disp_right->data && (!img_ok(disp_right, DWLS_16S, 1, 1) || disp_right->rows != h || disp_right->cols != w))
        return fail(ctx, DWLS_BAD_ARG, "disp_right must be CV_16SC1 of the view size");
    const bool both = disp_right && disp_right->data;
    CK(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    CK(call_enter(ctx, st));
    const size_t b_v = align256((size_t)w * h * cn), b_d = align256((size_t)w * h * 2);
    if ((s = ensure(ctx, &ctx->bmstage, &ctx->bmstage_bytes, 2 * b_v + 2 * b_d))) return s;
    uint8_t* d_l = (uint8_t*)ctx->bmstage; uint8_t* d_r = d_l + b_v; uint8_t* d_dl = d_r + b_v; uint8_t* d_dr = d_dl + b_d;
    CK(cudaMemcpy2DAsync(d_l, (size_t)w * cn, left->data, (size_t)left->step, (size_t)w * cn, h, cudaMemcpyHostToDevice, st));
    CK(cudaMemcpy2DAsync(d_r, (size_t)w * cn, right->data, (size_t)right->step, (size_t)w * cn, h, cudaMemcpyHostToDevice, st));
    ctx->h2d_bytes += 2 * (int64_t)w * h * cn;
    if ((s = bm_frames(ctx, d_l, (int64_t)w * cn, 0, d_r, (int64_t)w * cn, 0, cn, w, h, 1, p, (int16_t*)d_dl, w, 0, both ? (int16_t*)d_dr : nullptr, w, 0, st))) return s;
    CK(cudaMemcpy2DAsync(disp_left->data, (size_t)disp_left->step, d_dl, (size_t)w * 2, (size_t)w * 2, h, cudaMemcpyDeviceToHost, st));
    if (both) CK(cudaMemcpy2DAsync(disp_right->data, (size_t)disp_right->step, d_dr, (size_t)w * 2, (size_t)w * 2, h, cudaMemcpyDeviceToHost, st));
    ctx->d2h_bytes += (both ? 2 : 1) * (int64_t)w * h * 2;
    CK(cudaStreamSynchronize(st));
    return DWLS_OK;
}

// views halved like the sample's downscale step (cv::resize(..., Size(), 0.5, 0.5), samples/disparity_filtering.cpp:138-139):
// for even sizes bilinear interpolation at scale 0.5 is the rounded mean of each 2 x 2 block, (a + b + c + d + 2) >> 2
extern "C" dwls_status dwls_halve_views_dev(dwls_ctx* ctx, const dwls_views* v, uint8_t* left_out, uint8_t* right_out, int64_t out_step, int64_t out_frame, void* stream)
{
    if (!ctx) return DWLS_BAD_ARG;
    if (!v || v->n_frames <= 0 || !v->left || !v->right || !left_out || !right_out || (v->channels != 1 && v->channels != 3)) return fail(ctx, DWLS_BAD_ARG, "views / outputs missing");
    if ((v->cols & 1) || (v->rows & 1) || v->cols < 2 || v->rows < 2) return fail(ctx, DWLS_UNSUPPORTED, "the device path halves even-sized views only");
    CK(cudaSetDevice(ctx->device));
    cudaStream_t st = stream ? (cudaStream_t)stream : ctx->stream;
    CK(call_enter(ctx, st));
    HalfArgs ha;
    memset(&ha, 0, sizeof ha);
    ha.src[0] = v->left; ha.s_step[0] = v->left_step; ha.s_frame[0] = v->left_frame; ha.src[1] = v->right; ha.s_step[1] = v->right_step; ha.s_frame[1] = v->right_frame;
    ha.dst[0] = left_out; ha.dst[1] = right_out; ha.d_step = out_step; ha.d_frame = out_frame; ha.w = v->cols / 2; ha.h = v->rows / 2; ha.cn = v->channels;
    const int nb = ha.w * ha.cn;
    TIMED(K_HALF, st, half_kernel<<<dim3((nb + 255) / 256, ha.h, 2 * v->n_frames), 256, 0, st>>>(ha));
    CK(cudaGetLastError());
    CK(call_leave(ctx, st));
    return DWLS_OK;
}

// ---------------------------------------------------------------------------------------------
// fastGlobalSmootherFilter on device buffers WITHOUT any allocation (no counterpart in the reference; added
// for callers that cannot call cudaMalloc: code under CUDA-graph capture, tight serving
// loops, harnesses that forbid it).  Same kernels as dwls_fgs_dev; what that entry point keeps in cudaMalloc'ed memory
// (weight LUT, Chor, Cvert, up to 4 solver planes) lives here in ONE file-scope device array of a fixed, modest size.
// Every GPU that loads the library holds the array until the process exits.  All calls share it and it keeps the last
// call's contents: the plane region a call uses is cleared first, the LUT is rebuilt every call, and two calls that
// overlap on the GPU (different streams, no ordering between them) race on it - that is the caller's to prevent.
// ---------------------------------------------------------------------------------------------
static __device__ __align__(256) unsigned char g_noalloc_scratch[DWLS_NOALLOC_SCRATCH_BYTES];

// In host code `g_noalloc_scratch` is a host-side stand-in, not a device address: its device address is fetched once per
// device, with that device current, and that pointer is what the kernels get.
static std::mutex g_noalloc_mutex;
static unsigned char* g_noalloc_ptr[64] = {};

static dwls_status noalloc_scratch(dwls_ctx* ctx, unsigned char** out)
{
    if (ctx->device < 0 || ctx->device >= 64) return fail(ctx, DWLS_UNSUPPORTED, "device index beyond the no-allocation scratch table");
    std::lock_guard<std::mutex> lk(g_noalloc_mutex);
    if (!g_noalloc_ptr[ctx->device]) {
        void* p = nullptr;
        CK(cudaSetDevice(ctx->device));
        CK(cudaGetSymbolAddress(&p, g_noalloc_scratch));
        if (!p) return fail(ctx, DWLS_CUDA_ERROR, "cudaGetSymbolAddress returned a null scratch address");
        g_noalloc_ptr[ctx->device] = (unsigned char*)p;
    }
    *out = g_noalloc_ptr[ctx->device];
    return DWLS_OK;
}

extern "C" size_t dwls_fgs_noalloc_scratch_bytes(void) { return (size_t)DWLS_NOALLOC_SCRATCH_BYTES; }

// bytes of the static scratch a call on a w x h guide with `channels` source channels needs (0: not supported at all)
extern "C" size_t dwls_fgs_noalloc_required_bytes(int cols, int rows, int channels)
{
    if (cols <= 0 || rows <= 0 || channels < 1 || channels > 4) return 0;
    const size_t plane = (size_t)round_up(cols, 8) * rows;
    return align256(sizeof(float) * LUT_LEVELS) + (2 + (size_t)channels) * plane * sizeof(float);
}

extern "C" dwls_status dwls_fgs_dev_noalloc(dwls_ctx* ctx, const dwls_image* guide, const dwls_image* src, dwls_image* dst, double lambda,
                                            double sigma, double att, int num_iter, void* stream)
{
    if (!ctx) return DWLS_BAD_ARG;
    dwls_status s = fgs_check(ctx, guide, src, dst, lambda, sigma, num_iter);
    if (s) return s;
    if (ctx->timing) return fail(ctx, DWLS_UNSUPPORTED, "dwls_fgs_dev_noalloc: per-kernel timing creates CUDA events; switch it off for this entry point");
    const int w = guide->cols, h = guide->rows, pitch = round_up(w, 8), S = src->channels;
    SolvePlan pl;
    if (!plan_solver(w, h, &pl) || pl.vT)
        return fail(ctx, DWLS_UNSUPPORTED, "dwls_fgs_dev_noalloc: images of more than 8192 columns or 2176 rows need scratch planes this entry point does not have");
    const size_t need = dwls_fgs_noalloc_required_bytes(w, h, S);
    if (need > (size_t)DWLS_NOALLOC_SCRATCH_BYTES) {
        char b_[256];
        snprintf(b_, sizeof b_, "dwls_fgs_dev_noalloc: a %d x %d image with %d channel(s) needs %zu bytes of scratch, the static array holds %zu "
                                "(use dwls_fgs_dev, which allocates)", w, h, S, need, (size_t)DWLS_NOALLOC_SCRATCH_BYTES);
        ctx->err = b_;
        return DWLS_UNSUPPORTED;
    }
    CK(cudaSetDevice(ctx->device));
    unsigned char* base = nullptr;
    if ((s = noalloc_scratch(ctx, &base))) return s;
    cudaStream_t st = stream ? (cudaStream_t)stream : ctx->stream;
    const size_t plane = (size_t)pitch * h;
    float* lut = (float*)base;
    float* chor = (float*)(base + align256(sizeof(float) * LUT_LEVELS));
    float* cvert = chor + plane;
    float* planes = cvert + plane;
    // the array keeps what the last call left there: start the weight and solver planes of this call from zero
    CK(cudaMemsetAsync(chor, 0, (2 + (size_t)S) * plane * sizeof(float), st));
    lut_kernel<<<(LUT_LEVELS + 255) / 256, 256, 0, st>>>(lut, LUT_LEVELS, (float)sigma);
    CK(cudaGetLastError());
    ctx->launches++;
    // launch_weights reads the LUT pointer from the context: point it at the scratch LUT for the duration of this call
    // (host-side state only; the context's own LUT in device memory is neither touched nor freed)
    struct LutSwap {
        dwls_ctx* c; float* saved;
        LutSwap(dwls_ctx* ctx_, float* l) : c(ctx_), saved(ctx_->lut) { c->lut = l; }
        ~LutSwap() { c->lut = saved; }
    } swap_(ctx, lut);
    if ((s = launch_weights(ctx, (const uint8_t*)guide->data, guide->step, 0, guide->channels, w, h, pitch, chor, cvert, plane, 1, st, pl.wstrip))) return s;
    const bool direct = S == 1 && src->depth == DWLS_32F && pitch == w && src->step == (int64_t)w * 4 && dst->step == (int64_t)w * 4 &&
                        (((uintptr_t)src->data | (uintptr_t)dst->data) & 31) == 0;
    if (!direct) {
        if (src->depth == DWLS_8U) launch_unpack<uint8_t>(src, planes, plane, pitch, st);
        else if (src->depth == DWLS_16S) launch_unpack<int16_t>(src, planes, plane, pitch, st);
        else launch_unpack<float>(src, planes, plane, pitch, st);
        CK(cudaGetLastError());
        ctx->launches++;
    }
    for (int c0 = 0; c0 < S; c0 += 2) {
        const int ns = (S - c0 >= 2) ? 2 : 1;
        SolveArgs a;
        memset(&a, 0, sizeof a);
        a.w_frame = plane; a.p_frame = plane; a.pitch = pitch; a.w = w; a.h = h;
        for (int k = 0; k < ns; ++k) { a.in[k] = planes + (size_t)(c0 + k) * plane; a.out[k] = planes + (size_t)(c0 + k) * plane; }
        if (direct) { a.in[0] = (const float*)src->data; a.out[0] = (float*)dst->data; }
        if ((s = run_sweeps(ctx, a, chor, cvert, ns, FIN_FLOAT, pl, 1, lambda, att, num_iter, st, nullptr))) return s;
    }
    if (direct) return DWLS_OK;
    if (dst->depth == DWLS_8U) launch_pack<uint8_t>(planes, plane, pitch, dst, st);
    else if (dst->depth == DWLS_16S) launch_pack<int16_t>(planes, plane, pitch, dst, st);
    else launch_pack<float>(planes, plane, pitch, dst, st);
    CK(cudaGetLastError());
    ctx->launches++;
    return DWLS_OK;
}

// ---------------------------------------------------------------------------------------------
// peer-memory gather (one process per GPU on one NVLink / NVSwitch box; no counterpart in the reference, whose path is
// single-process).  The gather root allocates the buffer of the whole batch with dwls_peer_alloc and hands its 64-byte CUDA
// IPC handle to the other ranks (any byte transport: the Python side broadcasts it with torch.distributed); they map it with
// dwls_peer_open and PUSH every finished chunk of CV_16S results straight to its final place with a peer-to-peer
// cudaMemcpyAsync on a copy stream - the copy engines move the bytes over NVLink, no SM of the sender or of the root is
// involved - ordered after the kernels that produced the chunk and overlapping the kernels of the next one.
// (Why not NCCL send / recv: measured at 8 GPUs, 7 peers sending to one root reach about 0.4 TB/s into it - a few channels,
// i.e. CTAs, per peer - which bounds the whole step)
// ---------------------------------------------------------------------------------------------
extern "C" dwls_status dwls_peer_alloc(dwls_ctx* ctx, size_t bytes, void** dev_ptr, unsigned char* handle64)
{
    if (!ctx || !dev_ptr || !handle64 || bytes == 0) return ctx ? fail(ctx, DWLS_BAD_ARG, "dwls_peer_alloc: bad argument") : DWLS_BAD_ARG;
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "CUDA IPC handle size");
    CK(cudaSetDevice(ctx->device));
    void* p = nullptr;
    CK(cudaMalloc(&p, bytes));
    cudaIpcMemHandle_t h;
    const cudaError_t e = cudaIpcGetMemHandle(&h, p);
    if (e != cudaSuccess) { cudaFree(p); ctx->err = std::string("cudaIpcGetMemHandle failed: ") + cudaGetErrorString(e); return DWLS_CUDA_ERROR; }
    memcpy(handle64, &h, 64);
    *dev_ptr = p;
    return DWLS_OK;
}

extern "C" dwls_status dwls_peer_free(dwls_ctx* ctx, void* dev_ptr)
{
    if (!ctx) return DWLS_BAD_ARG;
    CK(cudaSetDevice(ctx->device));
    CK(cudaDeviceSynchronize());
    CK(cudaFree(dev_ptr));
    return DWLS_OK;
}

extern "C" dwls_status dwls_peer_open(dwls_ctx* ctx, const unsigned char* handle64, void** mapped)
{
    if (!ctx || !handle64 || !mapped) return ctx ? fail(ctx, DWLS_BAD_ARG, "dwls_peer_open: bad argument") : DWLS_BAD_ARG;
    CK(cudaSetDevice(ctx->device));
    cudaIpcMemHandle_t h;
    memcpy(&h, handle64, 64);
    CK(cudaIpcOpenMemHandle(mapped, h, cudaIpcMemLazyEnablePeerAccess));
    return DWLS_OK;
}

extern "C" dwls_status dwls_peer_close(dwls_ctx* ctx, void* mapped)
{
    if (!ctx) return DWLS_BAD_ARG;
    CK(cudaSetDevice(ctx->device));
    CK(cudaDeviceSynchronize());
    CK(cudaIpcCloseMemHandle(mapped));
    return DWLS_OK;
}

// copy `bytes` from local device memory to mapped peer memory on the context's copy stream, after everything queued so far on
// `after_stream` (the stream the producing kernels were launched on); returns at once
extern "C" dwls_status dwls_peer_push(dwls_ctx* ctx, void* dst_mapped, const void* src, size_t bytes, void* after_stream)
{
    if (!ctx || !dst_mapped || !src) return ctx ? fail(ctx, DWLS_BAD_ARG, "dwls_peer_push: bad argument") : DWLS_BAD_ARG;
    if (bytes == 0) return DWLS_OK;
    CK(cudaSetDevice(ctx->device));
    if (!ctx->peer_stream) {
        CK(cudaStreamCreateWithFlags(&ctx->peer_stream, cudaStreamNonBlocking));
        CK(cudaEventCreateWithFlags(&ctx->peer_ev, cudaEventDisableTiming));
        CK(cudaEventCreateWithFlags(&ctx->peer_done, cudaEventDisableTiming));
    }
    cudaStream_t prod = after_stream ? (cudaStream_t)after_stream : ctx->stream;
    CK(cudaEventRecord(ctx->peer_ev, prod));
    CK(cudaStreamWaitEvent(ctx->peer_stream, ctx->peer_ev, 0));
    CK(cudaMemcpyAsync(dst_mapped, src, bytes, cudaMemcpyDefault, ctx->peer_stream));
    ctx->peer_bytes += (int64_t)bytes;
    return DWLS_OK;
}

// `stream` waits for every push issued so far
extern "C" dwls_status dwls_peer_flush(dwls_ctx* ctx, void* stream)
{
    if (!ctx) return DWLS_BAD_ARG;
    if (!ctx->peer_stream) return DWLS_OK;
    CK(cudaSetDevice(ctx->device));
    CK(cudaEventRecord(ctx->peer_done, ctx->peer_stream));
    CK(cudaStreamWaitEvent(stream ? (cudaStream_t)stream : ctx->stream, ctx->peer_done, 0));
    return DWLS_OK;
}

extern "C" int64_t dwls_peer_bytes(const dwls_ctx* ctx) { return ctx ? ctx->peer_bytes : 0; }
