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
// 8-bit image.  A CTA handles WT_PX pixels of one row: the bytes of rows i and i+1 are staged in
// shared memory with aligned 128-bit loads (the ROI view starts at an arbitrary byte), each
// thread then forms the squared colour distances of 4 consecutive pixels, gathers the LUT and
// writes two float4.  Columns w..pitch-1 are zero-filled.
// grid = (ceil(pitch / WT_PX), h, frames)
// ---------------------------------------------------------------------------------------------
constexpr int WT_T = 256, WT_PX = 4 * WT_T;

template <int CN>
__global__ void __launch_bounds__(WT_T) weights_kernel(const uint8_t* __restrict__ guide, int64_t gstep, int64_t gframe,
                                                       int w, int h, int pitch, const float* __restrict__ lut,
                                                       float* __restrict__ chor, float* __restrict__ cvert, int64_t wframe)
{
    constexpr int ROWB = ((WT_PX + 1) * CN + 15 + 15) / 16 * 16;       // staged bytes per row incl. misalignment slack
    __shared__ __align__(16) uint8_t sm[2][ROWB];
    const int tid = threadIdx.x, i = blockIdx.y, frame = blockIdx.z, c0 = blockIdx.x * WT_PX;
    const int npx = min(WT_PX + 1, w - c0);                             // pixels available from c0 on (incl. the right neighbour)
    int mis[2] = { 0, 0 };
    if (npx > 0) {
#pragma unroll
        for (int rr = 0; rr < 2; ++rr) {
            if (i + rr >= h) break;
            const uint8_t* rp = guide + (size_t)frame * gframe + (size_t)(i + rr) * gstep + (size_t)c0 * CN;
            mis[rr] = (int)((uintptr_t)rp & 15);
            const uint4* base = (const uint4*)(rp - mis[rr]);           // an aligned 16-byte chunk holding >= 1 valid byte never crosses a page
            const int nch = (mis[rr] + npx * CN + 15) >> 4;
            for (int k = tid; k < nch; k += WT_T) ((uint4*)sm[rr])[k] = __ldg(base + k);
        }
    }
    __syncthreads();
    const int j0 = c0 + 4 * tid;
    if (j0 >= pitch) return;
    float ho[4] = { 0.f, 0.f, 0.f, 0.f }, vo[4] = { 0.f, 0.f, 0.f, 0.f };
    if (j0 < w) {
        const uint8_t* r0 = sm[0] + mis[0] + 4 * tid * CN;
        const uint8_t* r1 = sm[1] + mis[1] + 4 * tid * CN;
        const bool vr = i + 1 < h;
        int px[5][CN];
#pragma unroll
        for (int q = 0; q < 5; ++q)
#pragma unroll
            for (int ch = 0; ch < CN; ++ch) px[q][ch] = (j0 + q < w) ? r0[q * CN + ch] : 0;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            if (j0 + q >= w) break;
            int dh = 0, dv = 0;
#pragma unroll
            for (int ch = 0; ch < CN; ++ch) {
                const int t = px[q][ch] - px[q + 1][ch];
                dh += t * t;
                if (vr) { const int u = px[q][ch] - (int)r1[q * CN + ch]; dv += u * u; }
            }
            ho[q] = (j0 + q + 1 < w) ? __ldg(lut + dh) : 0.f;           // Chor[i][w-1] = 0 (614)
            vo[q] = vr ? __ldg(lut + dv) : 0.f;                          // Cvert[h-1][j] = 0 (658-660)
        }
    }
    const size_t o = (size_t)frame * wframe + (size_t)i * pitch + j0;
    *(float4*)(chor + o) = make_float4(ho[0], ho[1], ho[2], ho[3]);
    *(float4*)(cvert + o) = make_float4(vo[0], vo[1], vo[2], vo[3]);
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
