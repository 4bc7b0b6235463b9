// dwls_stages.cuh — everything around the sweeps: weight LUT and weight planes, confidence map
// (box variance + left-right consistency), plane preparation, bilinear upscale, ROI fill and the
// type conversions of the standalone FGS entry point.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "dwls_passes.cuh"
#include "dwls_conf.cuh"
#include "dwls_conf_run.cuh"

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
// 8-bit image.  A CTA covers 4*blockDim.x columns x WT_RB rows.  The WT_RB+1 guide row segments
// it needs are fetched by the TMA engine as 1-D bulk copies (each from the 16-byte boundary below
// the segment, so any ROI offset and row step works; an aligned 16-byte chunk holding a valid
// byte never leaves the allocation) onto one mbarrier.  Every thread then walks down the rows
// with its 4 pixels (+ right neighbour) of the current row in registers: squared colour
// distances by VABSDIFF4 + IDP.4A, two LUT gathers per pixel, two float4 stores per row.
// Columns w..pitch-1 are zero-filled.  grid = (ceil(pitch / (4*blockDim.x)), ceil(h / WT_RB), frames)
// dynamic smem = (WT_RB+1) * rowpitch + 4 * WT_LUT + 8
// ---------------------------------------------------------------------------------------------
constexpr int WT_RB = 8;
constexpr int WT_LUT = 2048;   // head of the LUT kept in shared memory (squared distances below 2048 are almost all of a natural image)
__host__ __device__ inline int weights_row_pitch(int nt, int cn) { return ((4 * nt + 1) * cn + 15 + 15) / 16 * 16 + 16; }

// pixels 0..4 of a thread (channels in the low bytes, high bytes zero) from the staged row; off = byte offset of pixel 0
template <int CN>
__device__ __forceinline__ void wt_pixels(const uint8_t* row, int off, uint32_t (&P)[5])
{
    const uint32_t* wp = (const uint32_t*)(row + (off & ~3));
    const int bs = (off & 3) * 8;
    if (CN == 3) {
        const uint32_t w0 = wp[0], w1 = wp[1], w2 = wp[2], w3 = wp[3], w4 = wp[4];
        const uint32_t b0 = __funnelshift_r(w0, w1, bs), b1 = __funnelshift_r(w1, w2, bs), b2 = __funnelshift_r(w2, w3, bs),
                       b3 = __funnelshift_r(w3, w4, bs);
        P[0] = b0 & 0xffffffu;
        P[1] = __funnelshift_r(b0, b1, 24) & 0xffffffu;
        P[2] = __funnelshift_r(b1, b2, 16) & 0xffffffu;
        P[3] = b2 >> 8;
        P[4] = b3 & 0xffffffu;
    } else {
        const uint32_t w0 = wp[0], w1 = wp[1], w2 = wp[2];
        const uint32_t b0 = __funnelshift_r(w0, w1, bs), b1 = __funnelshift_r(w1, w2, bs);
        P[0] = b0 & 0xffu; P[1] = (b0 >> 8) & 0xffu; P[2] = (b0 >> 16) & 0xffu; P[3] = b0 >> 24; P[4] = b1 & 0xffu;
    }
}
__device__ __forceinline__ int wt_dist2(uint32_t a, uint32_t b)
{
    const uint32_t d = __vabsdiffu4(a, b);
    return (int)__dp4a(d, d, 0u);
}

#ifndef DWLS_WT_MINB
#define DWLS_WT_MINB 4
#endif
template <int CN>
__global__ void __launch_bounds__(256, DWLS_WT_MINB) weights_kernel(const uint8_t* __restrict__ guide, int64_t gstep, int64_t gframe,
                                                      int w, int h, int pitch, int rowpitch, const float* __restrict__ lut,
                                                      float* __restrict__ chor, float* __restrict__ cvert, int64_t wframe, int vstrip)
{
    extern __shared__ __align__(128) uint8_t wsm[];
    float* slut = (float*)(wsm + (size_t)(WT_RB + 1) * rowpitch);
    uint64_t* bar = (uint64_t*)(slut + WT_LUT);
    const int tid = threadIdx.x, frame = blockIdx.z, c0 = blockIdx.x * 4 * blockDim.x, i0 = blockIdx.y * WT_RB;
    const int nrows = min(WT_RB + 1, h - i0);                           // guide rows staged
    const int npx = min(4 * (int)blockDim.x + 1, w - c0);               // pixels available from c0 on (incl. the right neighbour)
    const uint8_t* seg0 = guide + (size_t)frame * gframe + (size_t)i0 * gstep + (size_t)c0 * CN;
    if (tid == 0) { mbar_init(bar, npx > 0 ? nrows + 1 : 1); mbar_fence_init(); }
    __syncthreads();
    if (npx > 0) {
        if (tid == 31) {
            mbar_arrive_expect_tx(bar, WT_LUT * (uint32_t)sizeof(float));
            bulk_g2s(slut, lut, WT_LUT * (uint32_t)sizeof(float), bar);
        }
        if (tid < nrows) {
            const uint8_t* rp = seg0 + (size_t)tid * gstep;
            const int mis = (int)((uintptr_t)rp & 15);
            const uint32_t bytes = (uint32_t)((mis + npx * CN + 15) & ~15);
            mbar_arrive_expect_tx(bar, bytes);
            bulk_g2s(wsm + (size_t)tid * rowpitch, rp - mis, bytes, bar);
        }
        mbar_wait(bar, 0);
    }
    const int j0 = c0 + 4 * tid;
    if (j0 >= pitch) return;
    const int rows_out = min(WT_RB, h - i0);
    float* oh = chor + (size_t)frame * wframe + (size_t)i0 * pitch + j0;
    // Cvert is stored as strips of `vstrip` columns, each strip contiguous ([strip][row][vstrip]), so that the column pass
    // fetches a whole strip with one bulk copy; vstrip == pitch is the plain row-major plane
    float* ov = cvert + (size_t)frame * wframe + (size_t)(j0 / vstrip) * ((size_t)h * vstrip) + (size_t)i0 * vstrip + (j0 % vstrip);
    if (j0 >= w) {
        for (int r = 0; r < rows_out; ++r) {
            *(float4*)(oh + (size_t)r * pitch) = make_float4(0.f, 0.f, 0.f, 0.f);
            *(float4*)(ov + (size_t)r * vstrip) = make_float4(0.f, 0.f, 0.f, 0.f);
        }
        return;
    }
    const int nh = min(4, w - 1 - j0), nv = min(4, w - j0);             // pixels of the run with a right neighbour / inside the image
    uint32_t P[5], N[5];
    wt_pixels<CN>(wsm, (int)((uintptr_t)seg0 & 15) + 4 * tid * CN, P);
#pragma unroll
    for (int q = 0; q < 5; ++q) N[q] = 0;
#pragma unroll 2
    for (int r = 0; r < rows_out; ++r) {
        const bool vr = i0 + r + 1 < h;                                  // Cvert[h-1][j] = 0 (658-660)
        if (vr) wt_pixels<CN>(wsm + (size_t)(r + 1) * rowpitch, (int)((uintptr_t)(seg0 + (size_t)(r + 1) * gstep) & 15) + 4 * tid * CN, N);
        float ho[4], vo[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int dh2 = wt_dist2(P[q], P[q + 1]), dv2 = wt_dist2(P[q], N[q]);
            ho[q] = (q < nh) ? (dh2 < WT_LUT ? slut[dh2] : __ldg(lut + dh2)) : 0.f;   // Chor[i][w-1] = 0 (614)
            vo[q] = (vr && q < nv) ? (dv2 < WT_LUT ? slut[dv2] : __ldg(lut + dv2)) : 0.f;
        }
        *(float4*)(oh + (size_t)r * pitch) = make_float4(ho[0], ho[1], ho[2], ho[3]);
        *(float4*)(ov + (size_t)r * vstrip) = make_float4(vo[0], vo[1], vo[2], vo[3]);
#pragma unroll
        for (int q = 0; q < 5; ++q) P[q] = N[q];
    }
}

// everything outside the ROI: filtered output = 16*(min_disp-1) with min_disp forced to 0
// (disparity_filters.cpp:149,254,284) and, when exported, confidence = 0 (187) / mask = 0.
// One launch covers only the border: top/bottom bands (full rows) and side bands (rows of the ROI,
// columns outside it).
struct BorderArgs {
    int16_t* out; int64_t out_step, out_frame;       // elements; may be NULL
    float* conf; int64_t conf_step, conf_frame;      // elements; may be NULL
    uint8_t* mask; int64_t mask_step;                // may be NULL (frame 0 only)
    int W, H, rx, ry, rw, rh;
};
// grid = (blocks, frames); a warp takes one row of a band at a time, its lanes stride over the columns
__global__ void __launch_bounds__(256) border_kernel(const BorderArgs a)
{
    const int full_rows = a.H - a.rh, side_cols = a.W - a.rw;
    const int tasks = full_rows + (side_cols > 0 ? a.rh : 0);
    const int frame = blockIdx.y, lane = threadIdx.x & 31;
    const int warps = gridDim.x * (blockDim.x >> 5);
    for (int t = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); t < tasks; t += warps) {
        const bool side = t >= full_rows;
        int i = side ? a.ry + (t - full_rows) : t;
        if (!side && i >= a.ry) i += a.rh;
        const int n = side ? side_cols : a.W;
        int16_t* o = a.out ? a.out + (size_t)frame * a.out_frame + (size_t)i * a.out_step : nullptr;
        float* c = a.conf ? a.conf + (size_t)frame * a.conf_frame + (size_t)i * a.conf_step : nullptr;
        uint8_t* m = (a.mask && frame == 0) ? a.mask + (size_t)i * a.mask_step : nullptr;
        for (int k = lane; k < n; k += 32) {
            const int j = (side && k >= a.rx) ? k + a.rw : k;
            if (o) o[j] = (int16_t)-16;
            if (c) c[j] = 0.f;
            if (m) m[j] = 0;
        }
    }
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
