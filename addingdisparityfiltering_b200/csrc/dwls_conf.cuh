// dwls_conf.cuh — confidence stage (disparity_filters.cpp:161-210, 306-341, 349-373): depth
// discontinuity (box variance) map of a view's ROI crop, and for the left view the left-right
// consistency check, the *255 scaling, the exported map and the two solver planes.
//
// Exactness: boxFilter / sqrBoxFilter of the reference are separable running sums over integers;
// here the window sum (int32) and the window sum of squares (int64) are formed exactly by a
// vertical sliding window per column followed by a horizontal sliding window per run of 8 pixels,
// then the reference's rounding sequence is applied: float(sum) * float(1/k^2);
// (double)sumsq * double(1/k^2) -> float; unfused float ops (ComputeDepthDisc_ParBody, 349-373).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace dwls {

struct ConfArgs {
    const int16_t* dl; int64_t dl_step, dl_frame;     // elements
    const int16_t* dr; int64_t dr_step, dr_frame;
    int dw, dh;                                       // disparity map size
    int rx, ry, rw, rh;                               // valid_disp_ROI (228-233)
    int rrx;                                          // mirrored right ROI x (202-203)
    int radius;
    float scale_f; double scale_d;                    // 1/k^2 as float (boxFilter) and double (sqrBoxFilter)
    float roll_off;                                   // 0.001f / resize_factor^2 (359)
    int thresh;                                       // (int)(resize_factor * LRC_thresh) (318)
    float* conf_r; int64_t cr_frame; int cr_pitch;    // right discontinuity map, ROI-cropped workspace, 0..1
    float* conf; int64_t conf_step, conf_frame;       // exported map (elements), full disparity size, 0..255; may be NULL
    float* plane_dc; float* plane_c; int64_t p_frame; int pitch;   // ROI-cropped solver planes; NULL in half-res mode
    int il;                                           // planes interleaved sector by sector in ONE array at plane_dc (SolveArgs::il): d*c is plane 0, c plane 1
    uint8_t* mask; int64_t mask_step;                 // optional LRC decision export (frame 0 only)
    int skip_c;                                       // the exported map doubles as plane c (the first row pass reads it there): do not write plane_c
    int skip_dc;                                      // the first row pass forms d*c itself from the map and the left disparities: do not write plane_dc
};

// offset of (frame, row y, column c, plane s) in a solver plane pair, c = first column of an aligned run of <= 8 (SolveArgs::il)
__device__ __forceinline__ size_t plane_off(int il, int64_t p_frame, int pitch, int frame, int y, int c, int s)
{
    return il ? (size_t)frame * (2 * p_frame) + (size_t)y * (2 * pitch) + (size_t)(c >> 3) * 16 + s * 8 + (c & 7)
              : (size_t)frame * p_frame + (size_t)y * pitch + c;
}

__device__ __forceinline__ int reflect101(int p, int len)
{
    if (len == 1) return 0;
    while (p < 0 || p >= len) p = p < 0 ? -p : 2 * (len - 1) - p;
    return p;
}

constexpr int CF_T = 256;                 // threads per CTA
constexpr int CF_TW = 224;                // output columns per CTA (28 runs of 8)
constexpr int CF_RT = 8;                  // rows per inner iteration (28 runs x 8 rows = 224 working threads in the horizontal phase)
constexpr int CF_ROWS = 32;               // output rows per CTA
constexpr int CF_MAXR = 64;               // largest radius
constexpr int CF_OWN = (CF_TW + 2 * CF_MAXR + CF_T - 1) / CF_T;   // columns a thread may own in the vertical phase (2)
constexpr int CF_NC = CF_TW + 2 * CF_MAXR;                        // haloed columns at the largest radius
constexpr int CF_NCP = CF_NC + CF_NC / 8 + 8;                     // skewed row length: column c lives at c + c/8 (runs of 8 start on distinct banks)

__device__ __forceinline__ int cf_idx(int c) { return c + (c >> 3); }

// grid = (ceil(rw / CF_TW), ceil(rh / CF_ROWS), frames), block = CF_T, dynamic smem = conf_smem_bytes(r)
// smem rows only need the haloed width of the actual radius
__host__ __device__ inline int conf_row_len(int r) { const int nc = CF_TW + 2 * r; return nc + nc / 8 + 8; }
__host__ __device__ inline size_t conf_smem_bytes(int r) { return (size_t)CF_RT * conf_row_len(r) * (sizeof(long long) + 2 * sizeof(int)); }

// LEFT = false: right view, writes conf_r over the right ROI.
// LEFT = true : left view + LRC (306-341) + *255 (209) + export + solver planes (286-290).
template <bool LEFT>
__global__ void __launch_bounds__(CF_T, 3) conf_kernel(const ConfArgs a)
{
    extern __shared__ __align__(16) unsigned char cf_smem[];
    const int r = a.radius, tid = threadIdx.x, frame = blockIdx.z;
    const int nc = CF_TW + 2 * r, rl = conf_row_len(r);
    long long* cq_s = (long long*)cf_smem;                 // [CF_RT][rl]
    int* cs_s = (int*)(cq_s + (size_t)CF_RT * rl);         // [CF_RT][rl]
    int* ct_s = cs_s + (size_t)CF_RT * rl;                 // [CF_RT][rl] centre values (left view only)
    const int tx0 = blockIdx.x * CF_TW;                    // first output column of the tile (crop coordinates)
    const int y0 = blockIdx.y * CF_ROWS, y1 = min(y0 + CF_ROWS, a.rh);
    const int16_t* src = (LEFT ? a.dl + (size_t)frame * a.dl_frame : a.dr + (size_t)frame * a.dr_frame) + (size_t)a.ry * (LEFT ? a.dl_step : a.dr_step)
                         + (LEFT ? a.rx : a.rrx);
    const int64_t sstep = LEFT ? a.dl_step : a.dr_step;

    // ---- vertical phase state: running column sums of the columns this thread owns ----
    const int16_t* col[CF_OWN];
    int cs[CF_OWN]; long long cq[CF_OWN];
    bool own[CF_OWN];
#pragma unroll
    for (int o = 0; o < CF_OWN; ++o) {
        const int c = tid + o * CF_T;
        own[o] = c < nc;
        col[o] = src + reflect101(tx0 + c - r, a.rw);      // border of the CROP (168-174)
        cs[o] = 0; cq[o] = 0;
        if (own[o])
            for (int d = -r; d < r; ++d) { const int v = col[o][(size_t)reflect101(y0 + d, a.rh) * sstep]; cs[o] += v; cq[o] += v * v; }
    }
    // ---- horizontal phase role: row rr of the iteration, run g of 8 output columns ----
    const int rr = tid / (CF_TW / 8), g = tid - rr * (CF_TW / 8);
    const bool hworker = tid < CF_RT * (CF_TW / 8);
    const int cx0 = tx0 + 8 * g;                           // first output column of the run (crop coordinates)

    for (int yb = y0; yb < y1; yb += CF_RT) {
#pragma unroll
        for (int o = 0; o < CF_OWN; ++o) {
            if (!own[o]) continue;
            int vin[CF_RT], vout[CF_RT], vct[CF_RT];
            if (yb - r >= 0 && yb + CF_RT - 1 + r < a.rh) {             // CTA-uniform: rows away from the crop border need no reflection
                const int16_t* pin = col[o] + (size_t)(yb + r) * sstep;
                const int16_t* pout = col[o] + (size_t)(yb - r) * sstep;
                const int16_t* pct = col[o] + (size_t)yb * sstep;
#pragma unroll
                for (int i = 0; i < CF_RT; ++i) {
                    vin[i] = pin[(size_t)i * sstep];
                    vout[i] = pout[(size_t)i * sstep];
                    if (LEFT) vct[i] = pct[(size_t)i * sstep];
                }
            } else {
#pragma unroll
                for (int i = 0; i < CF_RT; ++i) {
                    vin[i] = col[o][(size_t)reflect101(yb + i + r, a.rh) * sstep];
                    vout[i] = col[o][(size_t)reflect101(yb + i - r, a.rh) * sstep];
                    if (LEFT) vct[i] = col[o][(size_t)min(yb + i, a.rh - 1) * sstep];
                }
            }
            const int p = cf_idx(tid + o * CF_T);
#pragma unroll
            for (int i = 0; i < CF_RT; ++i) {
                cs[o] += vin[i]; cq[o] += vin[i] * vin[i];
                cs_s[i * rl + p] = cs[o]; cq_s[i * rl + p] = cq[o];
                if (LEFT) ct_s[i * rl + p] = vct[i];
                cs[o] -= vout[i]; cq[o] -= vout[i] * vout[i];
            }
        }
        __syncthreads();
        const int y = yb + rr;
        if (hworker && y < y1 && cx0 < a.rw) {
            const int* csr = cs_s + rr * rl; const long long* cqr = cq_s + rr * rl;
            // output column x (tile-local) covers haloed columns x .. x+2r
            int s = 0; long long q = 0;
            for (int d = 0; d <= 2 * r; ++d) { const int p = 9 * g + d + (d >> 3); s += csr[p]; q += cqr[p]; }          // cf_idx(8g + d)
            float cv[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const float mean = __fmul_rn((float)s, a.scale_f);
                const float meansq = (float)__dmul_rn((double)q, a.scale_d);
                const float variance = __fsub_rn(meansq, __fmul_rn(mean, mean));
                cv[i] = fmaxf(__fsub_rn(1.0f, __fmul_rn(a.roll_off, variance)), 0.0f);
                if (i < 7) {
                    const int e = i + 2 * r + 1, pa = 9 * g + e + (e >> 3), ps = 9 * g + i;                     // cf_idx(8g + e), cf_idx(8g + i)
                    s += csr[pa] - csr[ps]; q += cqr[pa] - cqr[ps];
                }
            }
            const int nvalid = min(8, a.rw - cx0);
            if (!LEFT) {
                float* O = a.conf_r + (size_t)frame * a.cr_frame + (size_t)y * a.cr_pitch + cx0;
                if (nvalid == 8) { *(float4*)O = make_float4(cv[0], cv[1], cv[2], cv[3]); *((float4*)O + 1) = make_float4(cv[4], cv[5], cv[6], cv[7]); }
                else
#pragma unroll
                    for (int i = 0; i < 8; ++i) if (i < nvalid) O[i] = cv[i];
            } else {
                const int gx0 = a.rx + cx0, gy = a.ry + y;
                const int* ctr = ct_s + rr * rl;
                int L[8], R[8];
                float cr[8], dc[8];
                bool hit[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    L[i] = ctr[9 * g + i + r + ((i + r) >> 3)];
                    const int ridx = gx0 + i - (L[i] >> 4);                                            // 331
                    hit[i] = i < nvalid && ridx >= a.rrx && ridx < a.rrx + a.rw;                       // 332
                    R[i] = hit[i] ? a.dr[(size_t)frame * a.dr_frame + (size_t)gy * a.dr_step + ridx] : 0;
                    cr[i] = hit[i] ? a.conf_r[(size_t)frame * a.cr_frame + (size_t)y * a.cr_pitch + (ridx - a.rrx)] : 0.f;
                }
                uint8_t m[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    m[i] = 1;
                    if (hit[i]) {
                        if (abs(L[i] + R[i]) < a.thresh) { cv[i] = fminf(cv[i], cr[i]); m[i] = 2; }     // 334-335
                        else { cv[i] = 0.0f; m[i] = 3; }                                                // 337
                    }
                    cv[i] = __fmul_rn(255.0f, cv[i]);                                                  // 209
                    dc[i] = __fmul_rn(cv[i], (float)L[i]);                                             // 289-290
                }
                if (a.mask && frame == 0)
#pragma unroll
                    for (int i = 0; i < 8; ++i) if (i < nvalid) a.mask[(size_t)gy * a.mask_step + gx0 + i] = m[i];
                if (a.plane_c) {
                    float* PC = (a.il ? a.plane_dc : a.plane_c) + plane_off(a.il, a.p_frame, a.pitch, frame, y, cx0, 1);
                    float* PD = a.plane_dc + plane_off(a.il, a.p_frame, a.pitch, frame, y, cx0, 0);
                    if (nvalid == 8) {
                        if (!a.skip_c) { *(float4*)PC = make_float4(cv[0], cv[1], cv[2], cv[3]); *((float4*)PC + 1) = make_float4(cv[4], cv[5], cv[6], cv[7]); }
                        if (!a.skip_dc) { *(float4*)PD = make_float4(dc[0], dc[1], dc[2], dc[3]); *((float4*)PD + 1) = make_float4(dc[4], dc[5], dc[6], dc[7]); }
                    } else
#pragma unroll
                        for (int i = 0; i < 8; ++i) if (i < nvalid) { if (!a.skip_c) PC[i] = cv[i]; if (!a.skip_dc) PD[i] = dc[i]; }
                }
                if (a.conf) {
                    float* E = a.conf + (size_t)frame * a.conf_frame + (size_t)gy * a.conf_step + gx0;
                    if (nvalid == 8 && (((uintptr_t)E) & 15) == 0) {
                        *(float4*)E = make_float4(cv[0], cv[1], cv[2], cv[3]); *((float4*)E + 1) = make_float4(cv[4], cv[5], cv[6], cv[7]);
                    } else
#pragma unroll
                        for (int i = 0; i < 8; ++i) if (i < nvalid) E[i] = cv[i];
                }
            }
        }
        __syncthreads();                        // the shared rows are rewritten by the next iteration
    }
}

}  // namespace dwls
