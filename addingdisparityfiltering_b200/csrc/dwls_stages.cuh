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

constexpr int CT_W = 32, CT_H = 16;     // output tile of the confidence kernels

// discontinuity confidence of one view's ROI crop: max(1 - roll_off * (E[d^2] - E[d]^2), 0) with
// the exact rounding sequence of boxFilter / sqrBoxFilter / ComputeDepthDisc_ParBody:
// int window sum * float(1/k^2); exact sum of squares * double(1/k^2) -> float; unfused float ops.
__device__ __forceinline__ float disc_conf(const int16_t* tile, int tw, int lx, int ly, int r, const ConfArgs& a)
{
    int s = 0; long long q = 0;
    for (int dy = 0; dy <= 2 * r; ++dy) {
        const int16_t* t = tile + (ly + dy) * tw + lx;
        int rs = 0;                      // squares go straight into the 64-bit accumulator (a row of them can exceed 2^32)
        for (int dx = 0; dx <= 2 * r; ++dx) {
            const int v = t[dx];
            rs += v;
            q += (long long)(v * v);
        }
        s += rs;
    }
    const float mean = __fmul_rn((float)s, a.scale_f);
    const float meansq = (float)__dmul_rn((double)q, a.scale_d);
    const float variance = __fsub_rn(meansq, __fmul_rn(mean, mean));
    return fmaxf(__fsub_rn(1.0f, __fmul_rn(a.roll_off, variance)), 0.0f);
}

// grid = (ceil(dw/CT_W), ceil(dh/CT_H), frames), block = (CT_W, CT_H); dynamic smem = tile
// LEFT = false: right view, writes conf_r over the right ROI.
// LEFT = true : left view + LRC + *255 + export + planes.  Covers the whole disparity map so the
//               exported confidence is zero outside the ROI (187).
template <bool LEFT>
__global__ void __launch_bounds__(CT_W * CT_H) conf_kernel(const ConfArgs a)
{
    extern __shared__ int16_t tile[];
    const int r = a.radius, tw = CT_W + 2 * r, th = CT_H + 2 * r;
    const int frame = blockIdx.z;
    const int x0 = blockIdx.x * CT_W, y0 = blockIdx.y * CT_H;          // full-map coordinates (LEFT) / crop coordinates (!LEFT)
    const int16_t* src = LEFT ? a.dl + (size_t)frame * a.dl_frame : a.dr + (size_t)frame * a.dr_frame;
    const int64_t sstep = LEFT ? a.dl_step : a.dr_step;
    const int ox = LEFT ? a.rx : a.rrx;                                  // crop origin in the source map
    // tile origin in crop coordinates
    const int cx0 = LEFT ? x0 - a.rx : x0, cy0 = LEFT ? y0 - a.ry : y0;
    const bool tile_hits_roi = cx0 < a.rw && cx0 + CT_W > 0 && cy0 < a.rh && cy0 + CT_H > 0;
    if (tile_hits_roi) {
        for (int idx = threadIdx.y * CT_W + threadIdx.x; idx < tw * th; idx += CT_W * CT_H) {
            const int ty = idx / tw, tx = idx - ty * tw;
            const int cx = reflect101(cx0 + tx - r, a.rw), cy = reflect101(cy0 + ty - r, a.rh);   // border of the CROP (168-174)
            tile[idx] = src[(size_t)(a.ry + cy) * sstep + ox + cx];
        }
    }
    __syncthreads();
    const int cx = cx0 + threadIdx.x, cy = cy0 + threadIdx.y;
    const bool in_roi = cx >= 0 && cx < a.rw && cy >= 0 && cy < a.rh;
    if (!LEFT) {
        if (in_roi) a.conf_r[(size_t)frame * a.cr_frame + (size_t)cy * a.cr_pitch + cx] = disc_conf(tile, tw, threadIdx.x, threadIdx.y, r, a);
        return;
    }
    const int gx = x0 + threadIdx.x, gy = y0 + threadIdx.y;
    if (gx >= a.dw || gy >= a.dh) return;
    float cval = 0.f;
    if (in_roi) {
        cval = disc_conf(tile, tw, threadIdx.x, threadIdx.y, r, a);
        const int L = tile[(threadIdx.y + r) * tw + threadIdx.x + r];
        const int ridx = gx - (L >> 4);                                                       // 331
        uint8_t m = 1;
        if (ridx >= a.rrx && ridx < a.rrx + a.rw) {                                            // 332
            const int R = a.dr[(size_t)frame * a.dr_frame + (size_t)gy * a.dr_step + ridx];
            if (abs(L + R) < a.thresh) {                                                       // 334
                cval = fminf(cval, a.conf_r[(size_t)frame * a.cr_frame + (size_t)cy * a.cr_pitch + (ridx - a.rrx)]);
                m = 2;
            } else { cval = 0.0f; m = 3; }
        }
        cval = __fmul_rn(255.0f, cval);                                                        // 209
        if (a.mask && frame == 0) a.mask[(size_t)gy * a.mask_step + gx] = m;
        if (a.plane_c) {
            const size_t po = (size_t)frame * a.p_frame + (size_t)cy * a.pitch + cx;
            a.plane_c[po] = cval;
            a.plane_dc[po] = __fmul_rn(cval, (float)L);                                        // 289-290
        }
    } else if (a.mask && frame == 0) a.mask[(size_t)gy * a.mask_step + gx] = 0;
    if (a.conf) a.conf[(size_t)frame * a.conf_frame + (size_t)gy * a.conf_step + gx] = cval;
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

// output outside the ROI = 16*(min_disp-1) with min_disp forced to 0 (disparity_filters.cpp:149,254,284)
// grid = (ceil(W/256), H, frames)
__global__ void fill_outside_kernel(int16_t* __restrict__ out, int64_t ostep, int64_t oframe, int W, int H,
                                    int rx, int ry, int rw, int rh)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x, i = blockIdx.y, frame = blockIdx.z;
    if (j >= W) return;
    if (i >= ry && i < ry + rh && j >= rx && j < rx + rw) return;
    out[(size_t)frame * oframe + (size_t)i * ostep + j] = (int16_t)-16;
}

}  // namespace dwls
