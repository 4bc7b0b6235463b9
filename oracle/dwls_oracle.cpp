// dwls_oracle.cpp — CPU ORACLE (test infrastructure, NOT product code).
//
// A scalar/threaded CPU restatement of the arithmetic of the reference hot path
//   cv::ximgproc::DisparityWLSFilter::filter  (modules/ximgproc/src/disparity_filters.cpp:219-298)
//   cv::ximgproc::FastGlobalSmootherFilter    (modules/ximgproc/src/fgs_filter.cpp:141-233)
// on raw pointer+stride buffers.  Only tests/, __graft_entry__.smoke() and the
// cpu_baseline / --impl reference legs of bench.py may load this file's library.
// The product (addingdisparityfiltering_b200/) never links or calls it.
//
// Pinning status: the FGS solver, LRC, confidence formula and orchestration are pinned
// against the reference's own sources compiled from /root/reference (oracle/_ref, built by
// oracle/Makefile against the minimal OpenCV-core stand-in in oracle/shim/) — see
// tests/test_oracle_vs_ref.py and tests/golden/.  The OpenCV *imgproc* primitives the
// reference calls (boxFilter, sqrBoxFilter, resize, convertTo) are external to the reference
// tree; they are restated here from their published behaviour and cross-checked against
// cv2 4.13 in tests/test_oracle_primitives.py.
//
// Every function cites the reference file:line it follows.  Built with
// -ffp-contract=off so no FMA contraction changes the rounding sequence.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <functional>
#include <thread>
#include <vector>

namespace {

// ---------------------------------------------------------------------------------------------
// stripe-parallel helper: the reference fans every stage out with
// parallel_for_(Range(0,num_stripes), body) where stripe s covers
// [min(s*sz,n), min((s+1)*sz,n)), sz = ceil(n/num_stripes)   (fgs_filter.cpp:245-249, 466-470)
// ---------------------------------------------------------------------------------------------
struct Stripe { int begin, end; };

static inline Stripe stripe_of(int s, int nstripes, int n)
{
    int sz = (int)std::ceil(n / (double)nstripes);
    return { std::min(s * sz, n), std::min((s + 1) * sz, n) };
}

static void for_stripes(int nstripes, bool threaded, const std::function<void(int)>& body)
{
    if (!threaded || nstripes <= 1) { for (int s = 0; s < nstripes; ++s) body(s); return; }
    std::vector<std::thread> th;
    th.reserve(nstripes - 1);
    for (int s = 1; s < nstripes; ++s) th.emplace_back(body, s);
    body(0);
    for (auto& t : th) t.join();
}

// ---------------------------------------------------------------------------------------------
// numeric conversions OpenCV performs on this path (core: saturate_cast / cvRound):
// float -> int is round-half-to-even; NaN and |v| >= 2^31 give INT_MIN (x86 cvtss2si), which
// then saturates to the low end of the target type.
// ---------------------------------------------------------------------------------------------
static inline int cv_round(float v)
{
    if (!(std::fabs(v) < 2147483648.0f)) return INT32_MIN;
    return (int)std::nearbyintf(v);                 // default FE_TONEAREST = ties-to-even
}
static inline int16_t sat_s16(int v) { return (int16_t)std::min(std::max(v, -32768), 32767); }
static inline uint8_t sat_u8(int v) { return (uint8_t)std::min(std::max(v, 0), 255); }

// cv::borderInterpolate(p, len, BORDER_REFLECT_101)
static inline int reflect101(int p, int len)
{
    if (len == 1) return 0;
    while (p < 0 || p >= len) p = p < 0 ? -p : 2 * (len - 1) - p;
    return p;
}

// ---------------------------------------------------------------------------------------------
// FastGlobalSmoother
// ---------------------------------------------------------------------------------------------
enum Assoc { ASSOC_SCALAR = 0, ASSOC_SIMD_EMU = 1 };

struct Fgs
{
    int w = 0, h = 0, nstripes = 1;
    bool threaded = false;
    int assoc = ASSOC_SCALAR;
    float sigma = 1.f, lambda = 0.f, lambda_att = 0.25f;
    int num_iter = 3;
    std::vector<float> lut, chor, cvert, inter;

    // fgs_filter.cpp:141-173 (init) ; LUT at 669-675 ; weights at 593-616 and 625-661
    void init(const uint8_t* guide, int64_t gstep, int w_, int h_, int gcn, double lambda_, double sigma_,
              int num_iter_, double lambda_att_, int nstripes_, bool threaded_, int assoc_)
    {
        w = w_; h = h_; nstripes = std::max(1, nstripes_); threaded = threaded_; assoc = assoc_;
        sigma = (float)sigma_; lambda = (float)lambda_; lambda_att = (float)lambda_att_; num_iter = num_iter_;
        const int levels = 3 * 256 * 256;
        lut.resize(levels);
        for_stripes(nstripes, threaded, [&](int s) {
            Stripe r = stripe_of(s, nstripes, levels);
            for (int i = r.begin; i < r.end; ++i)
                lut[i] = (float)(-std::exp(-std::sqrt((float)i) / sigma));   // float exp/sqrt overloads (674)
        });
        chor.assign((size_t)w * h, 0.f); cvert.assign((size_t)w * h, 0.f); inter.assign((size_t)w * h, 0.f);
        auto dist = [&](const uint8_t* p, const uint8_t* q) {
            int d = 0;
            for (int c = 0; c < gcn; ++c) { int t = (int)p[c] - (int)q[c]; d += t * t; }   // 50-59
            return d;
        };
        // horizontal weights: Chor[i][j] = LUT[|g(i,j)-g(i,j+1)|^2], last column 0 (603-615)
        for_stripes(nstripes, threaded, [&](int s) {
            Stripe r = stripe_of(s, nstripes, h);
            for (int i = r.begin; i < r.end; ++i) {
                const uint8_t* g = guide + i * gstep;
                float* c = &chor[(size_t)i * w];
                for (int j = 0; j + 1 < w; ++j) c[j] = lut[dist(g + j * gcn, g + (j + 1) * gcn)];
                c[w - 1] = 0.f;
            }
        });
        // vertical weights: Cvert[i][j] = LUT[|g(i,j)-g(i+1,j)|^2], last row 0 (635-660)
        for_stripes(nstripes, threaded, [&](int s) {
            Stripe r = stripe_of(s, nstripes, w);
            for (int i = 0; i + 1 < h; ++i) {
                const uint8_t* g0 = guide + i * gstep; const uint8_t* g1 = guide + (i + 1) * gstep;
                float* c = &cvert[(size_t)i * w];
                for (int j = r.begin; j < r.end; ++j) c[j] = lut[dist(g0 + j * gcn, g1 + j * gcn)];
            }
            float* c = &cvert[(size_t)(h - 1) * w];
            for (int j = r.begin; j < r.end; ++j) c[j] = 0.f;
        });
    }

    // denominators: three association orders exist in the reference and any of them is
    // "the reference" depending on which rows/columns fall into SIMD blocks:
    static inline float den_scalar(float a, float c, float ip) { return (1 - a - c) - ip * a; }      // 369,455,553
    static inline float den_simd_h(float a, float c, float ip) { float t = ip * a; float u = c + a; u = 1.0f - u; return u - t; } // 306-310
    static inline float den_simd_v(float a, float c, float ip) { float t = ip * a; float u = a + c; u = u + t; return 1.0f - u; } // 530-534

    // one row, forward + backward (process_row, 439-464); simd_cols = number of leading columns
    // (beyond column 0) that a 4-row SIMD block would have covered (316: j<w-3 step 4).
    void solve_row(float* cur, const float* C, float* iD, float lam, bool in_4row_block) const
    {
        float a = lam * C[0];
        iD[0] = a / (1 - a);
        cur[0] = cur[0] / (1 - a);
        int simd_end = 1;
        if (in_4row_block && assoc == ASSOC_SIMD_EMU) { int j = 1; for (; j < w - 3; j += 4) {} simd_end = j; }
        for (int j = 1; j < w; ++j) {
            float c = lam * C[j];
            float den = (j < simd_end) ? den_simd_h(a, c, iD[j - 1]) : den_scalar(a, c, iD[j - 1]);
            iD[j] = c / den;
            cur[j] = (cur[j] - cur[j - 1] * a) / den;
            a = c;
        }
        for (int j = w - 2; j >= 0; --j) cur[j] = cur[j] - iD[j] * cur[j + 1];
    }

    // HorizontalPass_ParBody (466-476): inside a stripe rows go in groups of 4 (SIMD block), rest scalar
    void horizontal(float* cur, float lam)
    {
        for_stripes(nstripes, threaded, [&](int s) {
            Stripe r = stripe_of(s, nstripes, h);
            int i = r.begin;
            for (; i < r.end - 3; i += 4)
                for (int k = 0; k < 4; ++k) solve_row(cur + (size_t)(i + k) * w, &chor[(size_t)(i + k) * w], &inter[(size_t)(i + k) * w], lam, true);
            for (; i < r.end; ++i) solve_row(cur + (size_t)i * w, &chor[(size_t)i * w], &inter[(size_t)i * w], lam, false);
        });
    }

    // VerticalPass_ParBody (484-584): row-walking sweep over a stripe of columns
    void vertical(float* cur, float lam)
    {
        for_stripes(nstripes, threaded, [&](int s) {
            Stripe r = stripe_of(s, nstripes, w);
            int simd_end = r.begin;
            if (assoc == ASSOC_SIMD_EMU) simd_end = r.begin + 4 * ((r.end - r.begin) / 4);   // 520-521
            for (int j = r.begin; j < r.end; ++j) {                                           // 500-505
                float c = lam * cvert[j];
                inter[j] = c / (1 - c);
                cur[j] = cur[j] / (1 - c);
            }
            for (int i = 1; i < h; ++i) {                                                     // 506-557
                const float* Cp = &cvert[(size_t)(i - 1) * w]; const float* Cc = &cvert[(size_t)i * w];
                const float* iDp = &inter[(size_t)(i - 1) * w]; float* iD = &inter[(size_t)i * w];
                const float* up = cur + (size_t)(i - 1) * w; float* u = cur + (size_t)i * w;
                for (int j = r.begin; j < r.end; ++j) {
                    float a = lam * Cp[j], c = lam * Cc[j];
                    float den = (j < simd_end) ? den_simd_v(a, c, iDp[j]) : den_scalar(a, c, iDp[j]);
                    iD[j] = c / den;
                    u[j] = (u[j] - up[j] * a) / den;
                }
            }
            for (int i = h - 2; i >= 0; --i) {                                                // 560-583
                const float* iD = &inter[(size_t)i * w]; const float* un = cur + (size_t)(i + 1) * w; float* u = cur + (size_t)i * w;
                for (int j = r.begin; j < r.end; ++j) u[j] = u[j] - iD[j] * un[j];
            }
        });
    }

    // one float plane, in place: num_iter x (H, V, lambda *= att)   (fgs_filter.cpp:207-212)
    void filter_plane(float* cur)
    {
        float lam = lambda;
        for (int n = 0; n < num_iter; ++n) { horizontal(cur, lam); vertical(cur, lam); lam *= lambda_att; }
    }
};

} // namespace

// ---------------------------------------------------------------------------------------------
// exported C entry points (ctypes-friendly)
// ---------------------------------------------------------------------------------------------
extern "C" {

// depth codes follow OpenCV: 0 = CV_8U, 3 = CV_16S, 5 = CV_32F
enum { ORC_8U = 0, ORC_16S = 3, ORC_32F = 5 };

void orc_fgs_lut(float sigma, float* lut, int n)
{
    for (int i = 0; i < n; ++i) lut[i] = (float)(-std::exp(-std::sqrt((float)i) / sigma));
}

// weights only (for kernel-level parity of the weight planes)
int orc_fgs_weights(const uint8_t* guide, int64_t gstep, int w, int h, int gcn, double sigma, float* chor, float* cvert)
{
    if (!guide || w <= 0 || h <= 0 || (gcn != 1 && gcn != 3)) return 1;
    Fgs f; f.init(guide, gstep, w, h, gcn, 0.0, sigma, 1, 0.25, 1, false, ASSOC_SCALAR);
    std::memcpy(chor, f.chor.data(), sizeof(float) * (size_t)w * h);
    std::memcpy(cvert, f.cvert.data(), sizeof(float) * (size_t)w * h);
    return 0;
}

// fastGlobalSmootherFilter(guide, src, dst, lambda, sigma_color, lambda_attenuation, num_iter)
// (fgs_filter.cpp:687-691 -> 182-233): per channel convert to float, iterate, convert back.
int orc_fgs_filter(const uint8_t* guide, int64_t gstep, int w, int h, int gcn,
                   const void* src, int64_t sstep, int sdepth, int scn,
                   void* dst, int64_t dstep,
                   double lambda, double sigma, double lambda_att, int num_iter,
                   int assoc, int nstripes, int threaded)
{
    if (!guide || !src || !dst || w <= 0 || h <= 0) return 1;
    if ((gcn != 1 && gcn != 3) || lambda < 0 || sigma < 0 || num_iter < 1) return 1;            // 143-144
    if ((sdepth != ORC_8U && sdepth != ORC_16S && sdepth != ORC_32F) || scn < 1 || scn > 4) return 1; // 184
    Fgs f; f.init(guide, gstep, w, h, gcn, lambda, sigma, num_iter, lambda_att, nstripes, threaded != 0, assoc);
    std::vector<float> cur((size_t)w * h);
    for (int ch = 0; ch < scn; ++ch) {
        for (int i = 0; i < h; ++i) {
            const uint8_t* row = (const uint8_t*)src + i * sstep;
            float* c = &cur[(size_t)i * w];
            if (sdepth == ORC_8U) for (int j = 0; j < w; ++j) c[j] = (float)row[j * scn + ch];
            else if (sdepth == ORC_16S) for (int j = 0; j < w; ++j) c[j] = (float)((const int16_t*)row)[j * scn + ch];
            else for (int j = 0; j < w; ++j) c[j] = ((const float*)row)[j * scn + ch];
        }
        f.filter_plane(cur.data());
        for (int i = 0; i < h; ++i) {
            uint8_t* row = (uint8_t*)dst + i * dstep;
            const float* c = &cur[(size_t)i * w];
            if (sdepth == ORC_8U) for (int j = 0; j < w; ++j) row[j * scn + ch] = sat_u8(cv_round(c[j]));
            else if (sdepth == ORC_16S) for (int j = 0; j < w; ++j) ((int16_t*)row)[j * scn + ch] = sat_s16(cv_round(c[j]));
            else for (int j = 0; j < w; ++j) ((float*)row)[j * scn + ch] = c[j];
        }
    }
    return 0;
}

// ---- OpenCV imgproc primitives restated (external to the reference tree) ---------------------

// cv::boxFilter(src16S, dst32F, ksize k x k, anchor centre, normalize, BORDER_REFLECT_101):
// exact int32 window sum, then float(sum) * float(1/k^2)   (call site disparity_filters.cpp:108)
void orc_box_mean_16s(const int16_t* src, int64_t sstep_elems, int w, int h, int r, float* dst)
{
    const int k = 2 * r + 1; const float scale = (float)(1.0 / (k * k));
    std::vector<int> hs((size_t)w * h);
    for (int i = 0; i < h; ++i)
        for (int j = 0; j < w; ++j) { int s = 0; for (int d = -r; d <= r; ++d) s += src[i * sstep_elems + reflect101(j + d, w)]; hs[(size_t)i * w + j] = s; }
    for (int i = 0; i < h; ++i)
        for (int j = 0; j < w; ++j) { int s = 0; for (int d = -r; d <= r; ++d) s += hs[(size_t)reflect101(i + d, h) * w + j]; dst[(size_t)i * w + j] = (float)s * scale; }
}

// cv::sqrBoxFilter(src16S, dst32F, ...): double sum of squares, * (1/k^2) in double, cast to float
// (call site disparity_filters.cpp:114)
void orc_box_meansq_16s(const int16_t* src, int64_t sstep_elems, int w, int h, int r, float* dst)
{
    const int k = 2 * r + 1; const double scale = 1.0 / (k * k);
    std::vector<double> hs((size_t)w * h);
    for (int i = 0; i < h; ++i)
        for (int j = 0; j < w; ++j) { double s = 0; for (int d = -r; d <= r; ++d) { double v = src[i * sstep_elems + reflect101(j + d, w)]; s += v * v; } hs[(size_t)i * w + j] = s; }
    for (int i = 0; i < h; ++i)
        for (int j = 0; j < w; ++j) { double s = 0; for (int d = -r; d <= r; ++d) s += hs[(size_t)reflect101(i + d, h) * w + j]; dst[(size_t)i * w + j] = (float)(s * scale); }
}

} // extern "C" (templates below need C++ linkage)

// cv::resize(..., INTER_LINEAR) geometry: per destination index the source index and weight
static void linear_coeffs(int ssize, int dsize, std::vector<int>& ofs, std::vector<float>& frac, bool clamp_coeff)
{
    double inv_scale = (double)dsize / ssize; double scale = 1.0 / inv_scale;
    ofs.resize(dsize); frac.resize(dsize);
    for (int d = 0; d < dsize; ++d) {
        float f = (float)((d + 0.5) * scale - 0.5);
        int s = (int)std::floor(f);
        f -= s;
        if (clamp_coeff) {                 // horizontal: coefficient zeroed at the borders
            if (s < 0) { f = 0; s = 0; }
            if (s >= ssize - 1) { f = 0; s = ssize - 1; }
        }
        ofs[d] = s; frac[d] = f;
    }
}

// cv::resize bilinear, 1 channel, work type float, horizontal pass then vertical pass
// (call sites disparity_filters.cpp:243,272,274). T = int16_t (result RNE+saturate) or float.
template <typename T>
static void resize_linear(const T* src, int64_t sstep_elems, int sw, int sh, T* dst, int64_t dstep_elems, int dw, int dh)
{
    std::vector<int> xo, yo; std::vector<float> xf, yf;
    linear_coeffs(sw, dw, xo, xf, true);
    linear_coeffs(sh, dh, yo, yf, false);
    std::vector<float> hrow((size_t)sh * dw);
    for (int i = 0; i < sh; ++i) {
        const T* s = src + i * sstep_elems;
        for (int d = 0; d < dw; ++d) {
            int x = xo[d]; float a1 = xf[d], a0 = 1.f - a1;
            hrow[(size_t)i * dw + d] = (x >= sw - 1) ? (float)s[sw - 1] * 1.0f : (float)s[x] * a0 + (float)s[x + 1] * a1;
        }
    }
    for (int d = 0; d < dh; ++d) {
        int y0 = std::min(std::max(yo[d], 0), sh - 1), y1 = std::min(std::max(yo[d] + 1, 0), sh - 1);
        float b1 = yf[d], b0 = 1.f - b1;
        const float* S0 = &hrow[(size_t)y0 * dw]; const float* S1 = &hrow[(size_t)y1 * dw];
        T* o = dst + d * dstep_elems;
        for (int x = 0; x < dw; ++x) {
            float v = S0[x] * b0 + S1[x] * b1;
            if (sizeof(T) == 2) o[x] = (T)sat_s16(cv_round(v)); else o[x] = (T)v;
        }
    }
}

extern "C" {

void orc_resize_linear_16s(const int16_t* src, int64_t sstep_elems, int sw, int sh, int16_t* dst, int64_t dstep_elems, int dw, int dh)
{ resize_linear<int16_t>(src, sstep_elems, sw, sh, dst, dstep_elems, dw, dh); }
void orc_resize_linear_32f(const float* src, int64_t sstep_elems, int sw, int sh, float* dst, int64_t dstep_elems, int dw, int dh)
{ resize_linear<float>(src, sstep_elems, sw, sh, dst, dstep_elems, dw, dh); }

// ---- DisparityWLSFilter::filter ---------------------------------------------------------------

struct OrcWlsParams
{
    double lambda, sigma_color;          // disparity_filters.cpp:215 defaults 8000, 1.0
    int lrc_thresh, disc_radius;         // 154-155 defaults 24, 5
    int left_offset, right_offset, top_offset, bottom_offset;   // 144-145
    int use_confidence;
    int roi_x, roi_y, roi_w, roi_h;      // user ROI; area 0 => derive from offsets (228-233)
    int assoc, nstripes, threaded;       // oracle controls
};

// returns 0 ok, 1 bad argument (the reference would CV_Assert), 2 bad ROI
// out: filtered CV_16S at view size; conf_out (nullable): CV_32F confidence at view size when the
// disparity was upscaled else at disparity size (it is what getConfidenceMap() returns, 138, 274);
// roi_out: the disparity-resolution ROI getROI() returns (139); lrc_mask (nullable, dw x dh bytes):
// 0 outside ROI columns/rows, 1 right index outside right ROI (kept), 2 consistent, 3 inconsistent
int orc_wls_filter(const int16_t* dl, int64_t dlstep_elems, const int16_t* dr, int64_t drstep_elems, int dw, int dh,
                   const uint8_t* view, int64_t vstep, int W, int H, int vcn,
                   const OrcWlsParams* p,
                   int16_t* out, int64_t ostep_elems, float* conf_out, int64_t cstep_elems,
                   int* roi_out, uint8_t* lrc_mask)
{
    if (!dl || !view || !out || dw <= 0 || dh <= 0 || W <= 0 || H <= 0 || (vcn != 1 && vcn != 3)) return 1;   // 221-222
    const bool same_size = (dw == W && dh == H);
    const float resize_factor = same_size ? 1.0f : dw / (float)W;                                           // 224-227
    int rx, ry, rw, rh;
    if (p->roi_w * p->roi_h != 0) { rx = p->roi_x; ry = p->roi_y; rw = p->roi_w; rh = p->roi_h; }
    else { rx = p->left_offset; ry = p->top_offset; rw = dw - p->left_offset - p->right_offset; rh = dh - p->top_offset - p->bottom_offset; }
    if (roi_out) { roi_out[0] = rx; roi_out[1] = ry; roi_out[2] = rw; roi_out[3] = rh; }
    if (rx < 0 || ry < 0 || rw <= 0 || rh <= 0 || rx + rw > dw || ry + rh > dh) return 2;                  // Mat(m, roi) asserts
    const int nstripes = std::max(1, p->nstripes); const bool threaded = p->threaded != 0;

    std::vector<float> conf;            // disparity-resolution confidence, full dw x dh
    if (p->use_confidence) {
        if (!dr) return 1;                                                                                  // 262
        const int rrx = dw - (rx + rw);                                                                     // 202-203 mirrored ROI
        if (rrx < 0) return 2;
        const int r = p->disc_radius;
        std::vector<int16_t> cl((size_t)rw * rh), cr((size_t)rw * rh);                                      // copyToOp crops (168-174)
        for (int i = 0; i < rh; ++i) {
            std::memcpy(&cl[(size_t)i * rw], dl + (ry + i) * dlstep_elems + rx, sizeof(int16_t) * rw);
            std::memcpy(&cr[(size_t)i * rw], dr + (ry + i) * drstep_elems + rrx, sizeof(int16_t) * rw);
        }
        std::vector<float> ml((size_t)rw * rh), mr((size_t)rw * rh), ql((size_t)rw * rh), qr((size_t)rw * rh);
        for_stripes(4, threaded, [&](int s) {                                                               // 177-185
            if (s == 0) orc_box_mean_16s(cl.data(), rw, rw, rh, r, ml.data());
            if (s == 1) orc_box_mean_16s(cr.data(), rw, rw, rh, r, mr.data());
            if (s == 2) orc_box_meansq_16s(cl.data(), rw, rw, rh, r, ql.data());
            if (s == 3) orc_box_meansq_16s(cr.data(), rw, rw, rh, r, qr.data());
        });
        conf.assign((size_t)dw * dh, 0.f);                                                                  // 187
        std::vector<float> confr((size_t)dw * dh, 0.f);                                                     // 188
        const float roll_off = 0.001f / (resize_factor * resize_factor);                                    // 156, 359
        auto disc = [&](const std::vector<float>& m, const std::vector<float>& q, float* dstp, int x0) {    // 349-373
            for_stripes(nstripes, threaded, [&](int s) {
                Stripe st = stripe_of(s, nstripes, rh);
                for (int i = st.begin; i < st.end; ++i)
                    for (int j = 0; j < rw; ++j) {
                        float variance = q[(size_t)i * rw + j] - m[(size_t)i * rw + j] * m[(size_t)i * rw + j];
                        dstp[(size_t)(ry + i) * dw + x0 + j] = std::max(1.0f - roll_off * variance, 0.0f);
                    }
            });
        };
        disc(ml, ql, conf.data(), rx);
        disc(mr, qr, confr.data(), rrx);
        const int thresh = (int)(resize_factor * p->lrc_thresh);                                            // 318
        if (lrc_mask) std::memset(lrc_mask, 0, (size_t)dw * dh);
        for_stripes(nstripes, threaded, [&](int s) {                                                        // 306-341
            Stripe st = stripe_of(s, nstripes, dh);
            for (int i = st.begin; i < st.end; ++i) {
                const int16_t* L = dl + i * dlstep_elems; const int16_t* R = dr + i * drstep_elems;
                float* c = &conf[(size_t)i * dw]; const float* cR = &confr[(size_t)i * dw];
                const bool in_rows = (i >= ry && i < ry + rh);
                for (int j = rx; j < rx + rw; ++j) {
                    int ridx = j - (L[j] >> 4);
                    uint8_t m = 1;
                    if (ridx >= rrx && ridx < rrx + rw) {
                        if (std::abs((int)L[j] + (int)R[ridx]) < thresh) { c[j] = std::min(c[j], cR[ridx]); m = 2; }
                        else { c[j] = 0.0f; m = 3; }
                    }
                    if (lrc_mask && in_rows) lrc_mask[(size_t)i * dw + j] = m;
                }
            }
        });
        for (auto& v : conf) v = 255.0f * v;                                                                // 209
    }

    // optional upscale (239-247 / 268-277)
    std::vector<int16_t> disp_up; std::vector<float> conf_up;
    const int16_t* disp = dl; int64_t dstep = dlstep_elems;
    int fx = rx, fy = ry, fw = rw, fh = rh;
    if (!same_size) {
        float x_ratio = W / (float)dw, y_ratio = H / (float)dh;
        disp_up.resize((size_t)W * H);
        resize_linear<int16_t>(dl, dlstep_elems, dw, dh, disp_up.data(), W, W, H);
        for (auto& v : disp_up) v = sat_s16(cv_round((float)v * x_ratio));                                   // Mat*scalar -> convertTo, float work type
        disp = disp_up.data(); dstep = W;
        if (p->use_confidence) {
            conf_up.resize((size_t)W * H);
            resize_linear<float>(conf.data(), dw, dw, dh, conf_up.data(), W, W, H);
            conf.swap(conf_up);
        }
        fx = (int)(rx * x_ratio); fy = (int)(ry * y_ratio); fw = (int)(rw * x_ratio); fh = (int)(rh * y_ratio);
        if (fx < 0 || fy < 0 || fw <= 0 || fh <= 0 || fx + fw > W || fy + fh > H) return 2;
    }
    if (conf_out && p->use_confidence)
        for (int i = 0; i < H; ++i) std::memcpy(conf_out + i * cstep_elems, &conf[(size_t)i * W], sizeof(float) * W);

    for (int i = 0; i < H; ++i) for (int j = 0; j < W; ++j) out[i * ostep_elems + j] = (int16_t)(16 * (0 - 1));   // 149, 254/284

    Fgs f;
    f.init(view + (size_t)fy * vstep + (size_t)fx * vcn, vstep, fw, fh, vcn, p->lambda, p->sigma_color, 3, 0.25,
           nstripes, threaded, p->assoc);                                                                    // 257 / 292 defaults (edge_filter.hpp:393)
    std::vector<float> a((size_t)fw * fh);
    if (!p->use_confidence) {                                                                                // 235-259
        for (int i = 0; i < fh; ++i) for (int j = 0; j < fw; ++j) a[(size_t)i * fw + j] = (float)disp[(fy + i) * dstep + fx + j];
        f.filter_plane(a.data());
        for (int i = 0; i < fh; ++i) for (int j = 0; j < fw; ++j) out[(fy + i) * ostep_elems + fx + j] = sat_s16(cv_round(a[(size_t)i * fw + j]));
    } else {                                                                                                 // 286-296
        std::vector<float> c((size_t)fw * fh);
        for (int i = 0; i < fh; ++i) for (int j = 0; j < fw; ++j) {
            float cv = conf[(size_t)(fy + i) * W + fx + j];
            c[(size_t)i * fw + j] = cv;
            a[(size_t)i * fw + j] = cv * (float)disp[(fy + i) * dstep + fx + j];
        }
        f.filter_plane(a.data());
        f.filter_plane(c.data());
        const float eps = 1e-43f;                                                                            // 47
        for (int i = 0; i < fh; ++i) for (int j = 0; j < fw; ++j) {
            volatile float den = c[(size_t)i * fw + j] + eps;     // volatile: keep the denormal add under any -ffast flags
            float rec = (float)(1.0 / (double)den);                // cv::divide(1, M): scale/denom in double, cast to float
            float v = a[(size_t)i * fw + j] * rec;
            out[(fy + i) * ostep_elems + fx + j] = sat_s16(cv_round(v));
        }
    }
    return 0;
}

int orc_hardware_threads() { return (int)std::max(1u, std::thread::hardware_concurrency()); }

} // extern "C"
