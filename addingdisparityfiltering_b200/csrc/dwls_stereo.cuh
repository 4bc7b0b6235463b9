// dwls_stereo.cuh — block-matching stereo on the device (the step that PRODUCES the filter's inputs, so that
// the disparity maps are born in HBM instead of crossing PCIe).
//
// The reference takes its inputs from cv::StereoBM (OpenCV calib3d, outside /root/reference: the sample calls
// StereoBM::create(max_disp, wsize) + createRightMatcher, samples/disparity_filtering.cpp:148-160, under the parameter
// contract of src/disparity_filters.cpp:389-403, 419-428).  calib3d's source is not in the reference tree; the algorithm
// restated here is StereoBM's published one (x-Sobel prefilter with cap, SAD over a square window of the prefiltered
// views, winner-takes-all with optional texture / uniqueness tests, parabola sub-pixel fit, valid-disparity rectangle),
// pinned bit for bit against cv2.StereoBM.compute of the box's OpenCV 4.13 in tests/test_gpu_stereo.py (a numpy
// restatement checked against cv2 on the CPU lives with the other test infrastructure).
//
// Integer arithmetic throughout: the CV_16S result is bit-exact.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace dwls {

struct BmArgs {
    const uint8_t* left;  int64_t l_step, l_frame;       // prefiltered views (bytes)
    const uint8_t* right; int64_t r_step, r_frame;
    int16_t* disp; int64_t d_step, d_frame;               // elements
    int w, h;
    int ndisp, wsz, mindisp, ftzero, texture, uniq;
};

// ---------------------------------------------------------------------------------------------------------------
// prefilter: optional BGR -> gray (cv::cvtColor 8U: (3735 B + 19235 G + 9798 R + 2^14) >> 15), then StereoBM's
// PREFILTER_XSOBEL: rows are processed in pairs (y, y+1), y even, with the 1-2-1 column of x-differences taken over rows
// (y-1, y, y+1) and (y, y+1, y+2); row -1 mirrors to row 1 and a missing row y+2 to row y; the first and last column and,
// for an odd height, the last row hold the constant `cap`.  grid = (ceil(w/256), h, frames * views), block = 256
// ---------------------------------------------------------------------------------------------------------------
struct PrefArgs {
    const uint8_t* src[2]; int64_t s_step[2], s_frame[2]; int cn;      // two views (left, right), 1 or 3 channels
    uint8_t* dst[2]; int64_t d_step, d_frame;
    int w, h, cap;
};

__device__ __forceinline__ int bm_gray(const uint8_t* p, int cn)
{
    return cn == 1 ? (int)p[0] : (int)((3735u * p[0] + 19235u * p[1] + 9798u * p[2] + 16384u) >> 15);
}

__global__ void __launch_bounds__(256) bm_prefilter_kernel(const PrefArgs a)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y, v = blockIdx.z & 1, frame = blockIdx.z >> 1;
    if (x >= a.w) return;
    uint8_t* d = a.dst[v] + (size_t)frame * a.d_frame + (size_t)y * a.d_step + x;
    const bool paired = y < ((a.h - 1) & ~1) || ((a.h & 1) == 0);          // every row but the last one of an odd height
    if (x == 0 || x == a.w - 1 || !paired) { *d = (uint8_t)a.cap; return; }
    const uint8_t* s = a.src[v] + (size_t)frame * a.s_frame[v];
    const int ye = y & ~1;                                                   // the even row of this pair
    int r0, r1, r2;                                                          // rows weighted 1, 2, 1
    if (y == ye) { r1 = ye; r0 = ye > 0 ? ye - 1 : ye + 1; r2 = ye + 1; }
    else { r0 = ye; r1 = ye + 1; r2 = ye + 2 < a.h ? ye + 2 : ye; }
    auto dx = [&](int r) {
        const uint8_t* p = s + (size_t)r * a.s_step[v] + (size_t)x * a.cn;
        return bm_gray(p + a.cn, a.cn) - bm_gray(p - a.cn, a.cn);
    };
    const int val = dx(r0) + 2 * dx(r1) + dx(r2);
    *d = (uint8_t)(min(max(val, -a.cap), a.cap) + a.cap);
}

// ---------------------------------------------------------------------------------------------------------------
// SAD block matching + winner-takes-all + sub-pixel fit.
// A CTA owns BM_TW output columns x BM_RB rows.  Window column c of the tile (c = 0 .. BM_TW + wsz - 2) stands for the
// left column cl[c] and the right base column cr[c] (both clamped like StereoBM clamps them; candidate d reads right
// column cr[c] + d, i.e. disparity ndisp - 1 - d + mindisp).  For every row the CTA
//   A. updates the vertical window sums V[c][d] = sum over the wsz rows of |L - R| (row entering minus row leaving; the
//      sums live in registers, two 16-bit sums per register, four d per thread and step),
//   B. publishes them to shared memory as V2[d / 2][c] (column fastest),
//   C. forms SAD[x][d] = sum_{c = x .. x + wsz - 1} V[c][d] (thread = (x, quarter of the d range)): wsz CONSECUTIVE words
//      per pair of candidates, keeps the first minimum,
//   D. merges the four quarters, applies the texture / uniqueness tests and the parabola fit, and stores the CV_16S value.
// Pixels outside StereoBM's valid-disparity rectangle get FILTERED = (mindisp - 1) * 16.
// The kernel is bound by instruction issue (ncu, round 2: issue slots 68 % active, DRAM 0.06 %), so everything that does
// not depend on the row is computed once: the window size is a template parameter (WSZ = 0: any size, looped), and the
// (column, candidates) groups a thread owns are described by ONE packed register each, set up before the row loop
// (column | offset of the right view's bytes in a staged row | index of the pair of sums in V2).
// grid = (ceil(width1 / BM_TW), ceil(valid rows / BM_RB), frames * 2 views), block = 256
// ---------------------------------------------------------------------------------------------------------------
#ifndef DWLS_BM_RB
#define DWLS_BM_RB 32
#endif
constexpr int BM_TW = 64, BM_RB = DWLS_BM_RB, BM_T = 256;      // BM_RB: rows per band (each band re-does wsz - 1 warm-up rows)
constexpr int BM_MAXG = 16;      // (c, 4 d) groups a thread may own: (BM_TW + wsz - 1) * ndisp / 4 / BM_T <= BM_MAXG

struct BmPlan { int lofs, rofs, width1, y0, y1, xmin, xmax, wc, np, rw, lpitch; size_t smem; bool ok; };

__host__ __device__ inline BmPlan bm_plan(int w, int h, int ndisp, int wsz, int mindisp)
{
    BmPlan p;
    const int n = ndisp - 1 + mindisp, w2 = wsz / 2;
    p.lofs = n > 0 ? n : 0; p.rofs = n < 0 ? -n : 0;
    p.width1 = w - p.rofs - ndisp + 1;
    p.y0 = w2; p.y1 = h - w2;
    const int maxd = mindisp + ndisp - 1;
    p.xmin = (maxd > 0 ? maxd : 0) + w2;
    p.xmax = (w < w - mindisp ? w : w - mindisp) - w2;
    p.wc = BM_TW + wsz - 1;                     // window columns per tile
    p.np = ndisp / 2;                           // packed pairs of d
    p.lpitch = (p.wc + 3) & ~3;                 // staged left row (bytes)
    p.rw = (p.wc + ndisp + 3 + 3) & ~3;         // staged right row: wc + ndisp - 1 bytes (+ slack for word reads), multiple of 4
    const int rows = BM_RB + wsz - 1;
    p.smem = (size_t)rows * p.lpitch + (size_t)rows * p.rw + (size_t)p.np * p.wc * 4 + (size_t)p.wc * 4 + (size_t)BM_T * 8 + 64;
    // packed group descriptor: column < 256, right offset < 512, V2 index < 32768
    p.ok = p.lofs < w && p.rofs < w && p.width1 >= 1 && p.y1 > p.y0 && ndisp % 16 == 0 && ndisp >= 16 && (wsz & 1) && wsz >= 5 &&
           (p.wc * (ndisp / 4) + BM_T - 1) / BM_T <= BM_MAXG && p.wc < 256 && p.rw <= 512 && p.np * p.wc < 32768;
    return p;
}

// MAXG: groups a thread can own (8 or BM_MAXG = 16): the descriptors and running sums are registers, so the smaller capacity
// is used whenever the window columns x candidates of a tile fit it (e.g. 80 candidates, window 7: 6 groups per thread)
template <int WSZ, int MAXG>
__global__ void __launch_bounds__(BM_T) bm_kernel(const BmArgs a0, const BmArgs a1)
{
    extern __shared__ __align__(16) uint8_t bsm[];
    const int view = blockIdx.z & 1, frame = blockIdx.z >> 1;
    if (view == 1 && a1.disp == a0.disp) return;              // one matcher only: the second view's blocks have nothing to do
    const BmArgs& a = view ? a1 : a0;
    const int w = a.w, h = a.h, ndisp = a.ndisp, wsz = WSZ ? WSZ : a.wsz, w2 = wsz / 2;
    const BmPlan P = bm_plan(w, h, ndisp, wsz, a.mindisp);
    const int FILTERED = (a.mindisp - 1) * 16;
    const int tid = threadIdx.x;
    const int x0 = blockIdx.x * BM_TW;                        // relative output column (absolute = lofs + x)
    const int yb = P.y0 + blockIdx.y * BM_RB;                 // first output row of the band
    if (x0 >= P.width1 || yb >= P.y1) return;                 // (uniform) tiles of the other view's larger grid
    const int rows_out = min(BM_RB, P.y1 - yb);
    const int rows = rows_out + wsz - 1;                      // staged rows yb - w2 .. yb + rows_out - 1 + w2
    const int lpitch = P.lpitch, rw = P.rw, wc = P.wc;
    uint8_t* Ls = bsm;                                         // [rows][lpitch]
    uint8_t* Rs = Ls + (size_t)(BM_RB + wsz - 1) * lpitch;     // [rows][rw]
    uint32_t* V2 = (uint32_t*)(((uintptr_t)(Rs + (size_t)(BM_RB + wsz - 1) * rw) + 7) & ~(uintptr_t)7);   // [np][wc] packed pairs of 16-bit sums, column fastest (8-byte aligned: wc is even)
    int* T = (int*)(V2 + (size_t)P.np * wc);                   // [wc] vertical texture sums
    int* red = (int*)(((uintptr_t)(T + wc) + 7) & ~(uintptr_t)7);     // [2][BM_T] best key (SAD << 8 | d) per thread for its two pixels
    const uint8_t* Lg = a.left + (size_t)frame * a.l_frame;
    const uint8_t* Rg = a.right + (size_t)frame * a.r_frame;
    int16_t* Dg = a.disp + (size_t)frame * a.d_frame;

    // ---- stage the band: window column c <-> relative column xr = x0 - w2 + c ----
    const int xr0 = x0 - w2;
    auto colL = [&](int c) { return min(max(xr0 + c, -P.lofs), w - P.lofs - 1) + P.lofs; };
    auto colR = [&](int c) { return min(max(xr0 + c, -P.rofs), w - P.rofs - ndisp) + P.rofs; };
    const int rbase = colR(0);                                 // cr[] is non-decreasing: staged right columns rbase ..
    const int rcols = colR(wc - 1) + ndisp - rbase;            // .. rbase + rcols - 1  (<= wc + ndisp - 1)
    for (int idx = tid; idx < rows * wc; idx += BM_T) {
        const int r = idx / wc, c = idx - r * wc;
        Ls[r * lpitch + c] = Lg[(size_t)(yb - w2 + r) * a.l_step + colL(c)];
    }
    for (int idx = tid; idx < rows * rw; idx += BM_T) {
        const int r = idx / rw, j = idx - r * rw;
        Rs[r * rw + j] = j < rcols ? Rg[(size_t)(yb - w2 + r) * a.r_step + rbase + j] : (uint8_t)0;
    }

    // ---- the (window column, 4 candidates) groups of this thread, column fastest: g = d4 * wc + c ----
    const int nd4 = ndisp >> 2, ngroups = wc * nd4;
    uint32_t desc[MAXG];                                    // c | offR << 8 | vidx << 17
    int nk = 0;
#pragma unroll
    for (int k = 0; k < MAXG; ++k) {
        const int g = tid + k * BM_T;
        desc[k] = 0;
        if (g < ngroups) {
            const int d4 = g / wc, c = g - d4 * wc;
            const int offR = colR(c) - rbase + 4 * d4;         // bytes from the start of a staged right row
            desc[k] = (uint32_t)c | ((uint32_t)offR << 8) | ((uint32_t)(2 * d4 * wc + c) << 17);
            nk = k + 1;
        }
    }
    __syncthreads();

    uint32_t v01[MAXG], v23[MAXG];                       // 16-bit vertical sums of candidates d, d+1 | d+2, d+3
#pragma unroll
    for (int k = 0; k < MAXG; ++k) { v01[k] = 0; v23[k] = 0; }
    // |L - R| for the 4 candidates of a group in staged row r (Lrow / Rrow = that row's bytes; rows are word aligned)
    auto diff4 = [&](const uint8_t* Lrow, const uint8_t* Rrow, uint32_t dsc) {
        const uint32_t l4 = (uint32_t)Lrow[dsc & 0xffu] * 0x01010101u;
        const uint32_t off = (dsc >> 8) & 0x1ffu;
        const uint32_t* wp = (const uint32_t*)(Rrow + (off & ~3u));
        const uint32_t r4 = __funnelshift_r(wp[0], wp[1], (int)((off & 3u) * 8u));
        return __vabsdiffu4(l4, r4);
    };
    const bool want_tex = a.texture > 0;
    int tsum_c = 0;                                            // this thread's texture column (tid < wc)
    // warm-up: the first wsz - 1 rows of the window of output row yb
    for (int r = 0; r < wsz - 1; ++r) {
        const uint8_t* Lrow = Ls + r * lpitch; const uint8_t* Rrow = Rs + r * rw;
#pragma unroll
        for (int k = 0; k < MAXG; ++k)
            if (k < nk) {
                const uint32_t df = diff4(Lrow, Rrow, desc[k]);
                v01[k] += __byte_perm(df, 0, 0x4140); v23[k] += __byte_perm(df, 0, 0x4342);
            }
        if (want_tex && tid < wc) tsum_c += abs((int)Lrow[tid] - a.ftzero);
    }

    const int xq = tid & (BM_TW - 1), q = tid / BM_TW;         // phase C: output column of the tile, quarter of the d range
    const int dq0 = q * (ndisp / 4), dq1 = dq0 + ndisp / 4;    // candidates of this quarter (multiples of 4): uniqueness rescan
    const int xp = tid & 31, dg = tid >> 5;                    // phase C: pixel pair of the tile, eighth of the d range
    const int dg0 = dg * (ndisp >> 3), dg1 = dg0 + (ndisp >> 3);   // (multiples of 2: ndisp % 16 == 0)
    auto sad_at = [&](int x, int d) {                          // SAD of candidate d at tile column x from the published sums
        const uint32_t* p = V2 + (d >> 1) * wc + x;
        int s = 0;
        for (int k = 0; k < wsz; ++k) { const uint32_t v = p[k]; s += (d & 1) ? (int)(v >> 16) : (int)(v & 0xffffu); }
        return s;
    };

    for (int yo = 0; yo < rows_out; ++yo) {
        // A. row yo + wsz - 1 enters; row yo - 1 leaves (after the first output row)
        {
            const uint8_t* Lin = Ls + (yo + wsz - 1) * lpitch; const uint8_t* Rin = Rs + (yo + wsz - 1) * rw;
            const uint8_t* Lout = Ls + max(yo - 1, 0) * lpitch; const uint8_t* Rout = Rs + max(yo - 1, 0) * rw;
#pragma unroll
            for (int k = 0; k < MAXG; ++k)
                if (k < nk) {
                    const uint32_t dsc = desc[k];
                    const uint32_t dn = diff4(Lin, Rin, dsc);
                    v01[k] += __byte_perm(dn, 0, 0x4140); v23[k] += __byte_perm(dn, 0, 0x4342);
                    if (yo > 0) {
                        const uint32_t dl = diff4(Lout, Rout, dsc);
                        v01[k] -= __byte_perm(dl, 0, 0x4140); v23[k] -= __byte_perm(dl, 0, 0x4342);
                    }
                    // B. publish (sums stay below 2^16: wsz^2 * 2 * cap < 65536, checked on the host)
                    uint32_t* vp = V2 + (dsc >> 17);
                    vp[0] = v01[k]; vp[wc] = v23[k];
                }
            if (want_tex && tid < wc) {
                tsum_c += abs((int)Lin[tid] - a.ftzero);
                if (yo > 0) tsum_c -= abs((int)Lout[tid] - a.ftzero);
                T[tid] = tsum_c;
            }
        }
        __syncthreads();
        // C. thread (xp, dg): the pixel pair (2 xp, 2 xp + 1) and the dg-th eighth of the candidates.  The two window sums
        //    share all but one column, so one set of 64-bit loads (wsz + 1 consecutive words) serves both.  key = SAD << 8 | d:
        //    its integer minimum is the smallest SAD and, among equal SADs, the first candidate.
        int key0 = 0x7fffffff, key1 = 0x7fffffff;
        {
            const uint32_t* p = V2 + (dg0 >> 1) * wc + 2 * xp;
            for (int d = dg0; d < dg1; d += 2, p += wc) {
                uint32_t sa, sb;                                // both 16-bit halves at once (no carry: see B)
                if (WSZ) {
                    uint32_t wv[(WSZ ? WSZ : 1) + 1];
#pragma unroll
                    for (int k = 0; k < ((WSZ ? WSZ : 1) + 1) / 2; ++k) { const uint2 t2 = *(const uint2*)(p + 2 * k); wv[2 * k] = t2.x; wv[2 * k + 1] = t2.y; }
                    uint32_t mid = 0;
#pragma unroll
                    for (int k = 1; k < (WSZ ? WSZ : 1); ++k) mid += wv[k];
                    sa = mid + wv[0]; sb = mid + wv[WSZ ? WSZ : 1];
                } else {
                    sa = 0;
                    for (int k = 0; k < wsz; ++k) sa += p[k];
                    sb = sa - p[0] + p[wsz];                    // every half of sa contains the half of p[0]: no borrow
                }
                const int ka0 = (int)(((sa << 8) & 0x00ffff00u) | (uint32_t)d), ka1 = (int)(((sa >> 8) & 0x00ffff00u) | (uint32_t)(d + 1));
                const int kb0 = (int)(((sb << 8) & 0x00ffff00u) | (uint32_t)d), kb1 = (int)(((sb >> 8) & 0x00ffff00u) | (uint32_t)(d + 1));
                key0 = min(key0, min(ka0, ka1));
                key1 = min(key1, min(kb0, kb1));
            }
        }
        red[tid] = key0; red[BM_T + tid] = key1;
        __syncthreads();
        // D. merge the quarters (ascending d, strict <: the first minimum wins), tests, sub-pixel fit, store
        const int xrel = x0 + xq, xabs = P.lofs + xrel, y = yb + yo;
        bool filtered = false;
        int best = 0x7fffffff, bd = -1;
        auto merged_key = [&](int x) {                             // best key of tile pixel x over the eight candidate groups
            const int* r = red + (x & 1) * BM_T + (x >> 1);
            int m = r[0];
#pragma unroll
            for (int g8 = 1; g8 < 8; ++g8) m = min(m, r[g8 * 32]);
            return m;
        };
        if (q == 0 && xrel < P.width1) {
            const int m = merged_key(xq);
            best = m >> 8; bd = m & 0xff;
            if (want_tex) {
                int ts = 0;
                for (int c = xq; c < xq + wsz; ++c) ts += T[c];
                filtered = ts < a.texture;
            }
        }
        if (a.uniq > 0) {
            // every quarter rescans its candidates against the merged minimum (uniform branch: a.uniq is a kernel argument)
            const int mk = merged_key(xq);
            const int gb = mk >> 8, gd = mk & 0xff;
            const int thresh = gb + gb * a.uniq / 100;
            int hit = 0;
            for (int d = dq0; d < dq1; ++d)
                if ((d < gd - 1 || d > gd + 1) && sad_at(xq, d) <= thresh) { hit = 1; break; }
            __syncthreads();
            red[tid] = hit;
            __syncthreads();
            if (q == 0) filtered |= (red[xq] | red[xq + BM_TW] | red[xq + 2 * BM_TW] | red[xq + 3 * BM_TW]) != 0;
        }
        if (q == 0 && xrel < P.width1) {
            int val = FILTERED;
            if (!filtered && xabs >= P.xmin && xabs < P.xmax) {
                const int sp = sad_at(xq, bd + 1 < ndisp ? bd + 1 : ndisp - 2);         // sad[ndisp] = sad[ndisp - 2]
                const int sn = sad_at(xq, bd > 0 ? bd - 1 : 1);                           // sad[-1] = sad[1]
                const int dd = sp + sn - 2 * best + abs(sp - sn);
                val = ((ndisp - bd - 1 + a.mindisp) * 256 + (dd != 0 ? (sp - sn) * 256 / dd : 0) + 15) >> 4;
            }
            Dg[(size_t)y * a.d_step + xabs] = (int16_t)val;
        }
        __syncthreads();                                       // V2 / T / red are rewritten by the next row
    }
}

// everything the matcher does not compute: rows outside [y0, y1) and columns outside [lofs, lofs + width1)
// grid = (ceil(w/256), h, frames * 2), block = 256
__global__ void __launch_bounds__(256) bm_fill_kernel(const BmArgs a0, const BmArgs a1)
{
    const int view = blockIdx.z & 1, frame = blockIdx.z >> 1;
    if (view == 1 && a1.disp == a0.disp) return;
    const BmArgs& a = view ? a1 : a0;
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    if (x >= a.w || y >= a.h) return;
    const BmPlan P = bm_plan(a.w, a.h, a.ndisp, a.wsz, a.mindisp);
    const bool computed = P.lofs < a.w && P.rofs < a.w && P.width1 >= 1 && y >= P.y0 && y < P.y1 && x >= P.lofs && x < P.lofs + P.width1;
    if (!computed) a.disp[(size_t)frame * a.d_frame + (size_t)y * a.d_step + x] = (int16_t)((a.mindisp - 1) * 16);
}

// views halved: out(y, x, ch) = (s(2y, 2x) + s(2y, 2x+1) + s(2y+1, 2x) + s(2y+1, 2x+1) + 2) >> 2 per channel, which is what
// cv::resize(INTER_LINEAR) computes at scale 0.5 for even sizes.  grid = (ceil(w*cn/256), h, frames * 2), block = 256
struct HalfArgs {
    const uint8_t* src[2]; int64_t s_step[2], s_frame[2];
    uint8_t* dst[2]; int64_t d_step, d_frame;
    int w, h, cn;                                       // output size
};
__global__ void __launch_bounds__(256) half_kernel(const HalfArgs a)
{
    const int k = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y, v = blockIdx.z & 1, frame = blockIdx.z >> 1;
    if (k >= a.w * a.cn) return;
    const int x = k / a.cn, ch = k - x * a.cn;
    const uint8_t* s = a.src[v] + (size_t)frame * a.s_frame[v] + (size_t)(2 * y) * a.s_step[v] + (size_t)(2 * x) * a.cn + ch;
    const int sum = s[0] + s[a.cn] + s[a.s_step[v]] + s[a.s_step[v] + a.cn];
    a.dst[v][(size_t)frame * a.d_frame + (size_t)y * a.d_step + k] = (uint8_t)((sum + 2) >> 2);
}

}  // namespace dwls
