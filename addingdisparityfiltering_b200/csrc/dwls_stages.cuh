// dwls_stages.cuh — everything around the sweeps: weight LUT and weight planes, confidence map
// (box variance + left-right consistency), plane preparation, bilinear upscale, ROI fill and the
// type conversions of the standalone FGS entry point.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "dwls_passes.cuh"

namespace dwls {

// ---------------------------------------------------------------------------------------------
// LUT[i] = -exp(-sqrt(i)/sigma), i in [0, 3*256*256)   (fgs_filter.cpp:669-675)
// The reference evaluates exp in float on the host (glibc expf, <= 0.502 ulp).  Here: IEEE
// sqrt and divide in float, exp in double and one rounding to float, which is the correctly
// rounded float result except on double-rounding ties.
// ---------------------------------------------------------------------------------------------
__global__ void lut_kernel(float* lut, int n, float sigma)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float arg = -__fdiv_rn(__fsqrt_rn((float)i), sigma);
    lut[i] = -(float)exp((double)arg);
}

// ---------------------------------------------------------------------------------------------
// weight planes (fgs_filter.cpp:593-616 horizontal, 625-661 vertical), guide = ROI view of the
// 8-bit image.  One thread per pixel; columns w..pitch-1 are zero-filled.
// grid = (ceil(pitch/256), h, frames)
// ---------------------------------------------------------------------------------------------
template <int CN>
__global__ void __launch_bounds__(256) weights_kernel(const uint8_t* __restrict__ guide, int64_t gstep, int64_t gframe,
                                                      int w, int h, int pitch, const float* __restrict__ lut,
                                                      float* __restrict__ chor, float* __restrict__ cvert, int64_t wframe)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x, i = blockIdx.y, frame = blockIdx.z;
    if (j >= pitch) return;
    const size_t o = (size_t)frame * wframe + (size_t)i * pitch + j;
    if (j >= w) { chor[o] = 0.f; cvert[o] = 0.f; return; }
    const uint8_t* g = guide + (size_t)frame * gframe + (size_t)i * gstep + (size_t)j * CN;
    int dh = 0, dv = 0;
    const bool hr = j + 1 < w, vr = i + 1 < h;
#pragma unroll
    for (int ch = 0; ch < CN; ++ch) {
        const int p = g[ch];
        if (hr) { const int t = p - (int)g[CN + ch]; dh += t * t; }
        if (vr) { const int t = p - (int)g[gstep + ch]; dv += t * t; }
    }
    chor[o] = hr ? __ldg(lut + dh) : 0.f;
    cvert[o] = vr ? __ldg(lut + dv) : 0.f;
}

// ---------------------------------------------------------------------------------------------
// confidence stage (disparity_filters.cpp:161-210, 306-341, 349-373)
// ---------------------------------------------------------------------------------------------
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
    uint8_t* mask; int64_t mask_step;                 // optional LRC decision export (frame 0 only)
};

__device__ __forceinline__ int reflect101(int p, int len)
{
    if (len == 1) return 0;
    while (p < 0 || p >= len) p = p < 0 ? -p : 2 * (len - 1) - p;
    return p;
}

constexpr int CF_T = 256;       // threads per CTA = tile columns including the 2*radius halo
constexpr int CF_ROWS = 32;     // output rows per CTA
constexpr int CF_PF = 4;        // rows fetched ahead per thread

// Box variance by running sums, exact in integers: every thread owns one column of the ROI crop
// (border = reflect-101 of the CROP, disparity_filters.cpp:168-174) and slides a (2r+1)-row window
// down it (column sum, column sum of squares); the per-row column sums go through a double-buffered
// shared row and each output thread adds its 2r+1 neighbours.  Then the exact rounding sequence of
// boxFilter / sqrBoxFilter / ComputeDepthDisc_ParBody (349-373): int window sum * float(1/k^2);
// exact sum of squares * double(1/k^2) -> float; unfused float ops.
// grid = (ceil(rw / (CF_T - 2r)), ceil(rh / CF_ROWS), frames)
// LEFT = false: right view, writes conf_r over the right ROI.
// LEFT = true : left view + LRC (306-341) + *255 (209) + export + solver planes (286-290).
template <bool LEFT>
__global__ void __launch_bounds__(CF_T) conf_kernel(const ConfArgs a)
{
    __shared__ int cs_buf[2][CF_T];
    __shared__ long long cq_buf[2][CF_T];
    const int r = a.radius, tid = threadIdx.x, frame = blockIdx.z;
    const int tw = CF_T - 2 * r;
    const int cx = blockIdx.x * tw + tid - r;                      // crop column (halo threads fall outside [0, rw))
    const int y0 = blockIdx.y * CF_ROWS, y1 = min(y0 + CF_ROWS, a.rh);
    const int16_t* src = LEFT ? a.dl + (size_t)frame * a.dl_frame : a.dr + (size_t)frame * a.dr_frame;
    const int64_t sstep = LEFT ? a.dl_step : a.dr_step;
    const int16_t* col = src + (size_t)a.ry * sstep + (LEFT ? a.rx : a.rrx) + reflect101(cx, a.rw);
    const bool is_out = tid >= r && tid < CF_T - r && cx < a.rw;

    int cs = 0; long long cq = 0;
    for (int d = -r; d < r; ++d) { const int v = col[(size_t)reflect101(y0 + d, a.rh) * sstep]; cs += v; cq += v * v; }
    for (int yb = y0; yb < y1; yb += CF_PF) {
        int vin[CF_PF], vout[CF_PF];
#pragma unroll
        for (int i = 0; i < CF_PF; ++i) {
            vin[i] = col[(size_t)reflect101(yb + i + r, a.rh) * sstep];
            vout[i] = col[(size_t)reflect101(yb + i - r, a.rh) * sstep];
        }
#pragma unroll
        for (int i = 0; i < CF_PF; ++i) {
            const int y = yb + i;
            if (y >= y1) break;                                     // uniform across the CTA
            cs += vin[i]; cq += vin[i] * vin[i];
            const int buf = i & 1;                                  // CF_PF is even, so this alternates across chunks too
            cs_buf[buf][tid] = cs; cq_buf[buf][tid] = cq;
            __syncthreads();
            if (is_out) {
                int s = 0; long long q = 0;
                for (int d = -r; d <= r; ++d) { s += cs_buf[buf][tid + d]; q += cq_buf[buf][tid + d]; }
                const float mean = __fmul_rn((float)s, a.scale_f);
                const float meansq = (float)__dmul_rn((double)q, a.scale_d);
                const float variance = __fsub_rn(meansq, __fmul_rn(mean, mean));
                float cval = fmaxf(__fsub_rn(1.0f, __fmul_rn(a.roll_off, variance)), 0.0f);
                if (!LEFT) {
                    a.conf_r[(size_t)frame * a.cr_frame + (size_t)y * a.cr_pitch + cx] = cval;
                } else {
                    const int gx = a.rx + cx, gy = a.ry + y;
                    const int L = col[(size_t)y * sstep];
                    const int ridx = gx - (L >> 4);                                                   // 331
                    uint8_t m = 1;
                    if (ridx >= a.rrx && ridx < a.rrx + a.rw) {                                        // 332
                        const int R = a.dr[(size_t)frame * a.dr_frame + (size_t)gy * a.dr_step + ridx];
                        if (abs(L + R) < a.thresh) {                                                   // 334
                            cval = fminf(cval, a.conf_r[(size_t)frame * a.cr_frame + (size_t)y * a.cr_pitch + (ridx - a.rrx)]);
                            m = 2;
                        } else { cval = 0.0f; m = 3; }
                    }
                    cval = __fmul_rn(255.0f, cval);                                                    // 209
                    if (a.mask && frame == 0) a.mask[(size_t)gy * a.mask_step + gx] = m;
                    if (a.plane_c) {
                        const size_t po = (size_t)frame * a.p_frame + (size_t)y * a.pitch + cx;
                        a.plane_c[po] = cval;
                        a.plane_dc[po] = __fmul_rn(cval, (float)L);                                    // 289-290
                    }
                    if (a.conf) a.conf[(size_t)frame * a.conf_frame + (size_t)gy * a.conf_step + gx] = cval;
                }
            }
            cs -= vout[i]; cq -= vout[i] * vout[i];
        }
    }
}

// everything outside the ROI: filtered output = 16*(min_disp-1) with min_disp forced to 0
// (disparity_filters.cpp:149,254,284) and, when exported, confidence = 0 (187) / mask = 0.
// Two launches cover only the border: side bands (rows of the ROI, columns outside it) and the
// top/bottom bands (full rows).  grid = (ceil(cols/256), rows, frames)
struct BorderArgs {
    int16_t* out; int64_t out_step, out_frame;       // elements; may be NULL
    float* conf; int64_t conf_step, conf_frame;      // elements; may be NULL
    uint8_t* mask; int64_t mask_step;                // may be NULL (frame 0 only)
    int W, H, rx, ry, rw, rh;
};
template <bool SIDES>
__global__ void __launch_bounds__(256) border_kernel(const BorderArgs a)
{
    int j = blockIdx.x * blockDim.x + threadIdx.x, i = blockIdx.y;
    const int frame = blockIdx.z;
    if (SIDES) {
        if (j >= a.W - a.rw) return;
        if (j >= a.rx) j += a.rw;
        i += a.ry;
    } else {
        if (j >= a.W) return;
        if (i >= a.ry) i += a.rh;
    }
    if (a.out) a.out[(size_t)frame * a.out_frame + (size_t)i * a.out_step + j] = (int16_t)-16;
    if (a.conf) a.conf[(size_t)frame * a.conf_frame + (size_t)i * a.conf_step + j] = 0.f;
    if (a.mask && frame == 0) a.mask[(size_t)i * a.mask_step + j] = 0;
}

// ---------------------------------------------------------------------------------------------
// no-confidence path: plane = float(disparity ROI)   (disparity_filters.cpp:250, fgs 203-205)
// grid = (ceil(w/256), h, frames)
// ---------------------------------------------------------------------------------------------
__global__ void prep_noconf_kernel(const int16_t* __restrict__ disp, int64_t dstep, int64_t dframe,
                                   int w, int h, float* __restrict__ plane, int64_t pframe, int pitch)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x, i = blockIdx.y, frame = blockIdx.z;
    if (j >= w) return;
    plane[(size_t)frame * pframe + (size_t)i * pitch + j] = (float)disp[(size_t)frame * dframe + (size_t)i * dstep + j];
}

}  // namespace dwls
