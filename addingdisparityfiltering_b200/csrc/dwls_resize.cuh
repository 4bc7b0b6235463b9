// dwls_resize.cuh — half-resolution mode (disparity_filters.cpp:239-247, 268-277): bilinear
// upscale of the CV_16S disparity (x x_ratio, RNE + saturate) and of the CV_32F confidence to the
// view size, fused with plane preparation; plus the interleaved <-> planar conversions of the
// standalone fastGlobalSmootherFilter entry point (fgs_filter.cpp:191-205, 214-232).
#pragma once
#include "dwls_stages.cuh"

namespace dwls {

// cv::resize INTER_LINEAR geometry for one destination index (work type float):
// f = (float)((d+0.5)*scale-0.5); s = floor(f); f -= s.  Horizontal coefficients are zeroed at
// the borders, vertical source rows are clipped instead.
struct Lin { int s; float f; };
__device__ __forceinline__ Lin lin_coeff(int d, double scale, int ssize, bool clamp_coeff)
{
    float f = (float)(((double)d + 0.5) * scale - 0.5);
    int s = (int)floorf(f);
    f = __fsub_rn(f, (float)s);
    if (clamp_coeff) {
        if (s < 0) { f = 0.f; s = 0; }
        if (s >= ssize - 1) { f = 0.f; s = ssize - 1; }
    }
    return { s, f };
}

template <typename T>
__device__ __forceinline__ float lin_h(const T* row, Lin cx, int sw)
{
    if (cx.s >= sw - 1) return __fmul_rn((float)row[sw - 1], 1.0f);
    const float a1 = cx.f, a0 = __fsub_rn(1.f, a1);
    return __fadd_rn(__fmul_rn((float)row[cx.s], a0), __fmul_rn((float)row[cx.s + 1], a1));
}

struct UpArgs {
    const int16_t* disp; int64_t d_step, d_frame;     // low-res disparity (elements)
    const float* conf_lo; int64_t cl_step, cl_frame;  // low-res confidence 0..255 (NULL when !use_confidence)
    int dw, dh, W, H;
    double scale_x, scale_y;                          // 1 / (W/dw), 1 / (H/dh)
    float x_ratio;                                    // W / (float)dw
    int fx, fy, fw, fh;                               // ROI in view coordinates (245-246 / 275-276)
    float* conf; int64_t conf_step, conf_frame;       // exported view-size confidence (may be NULL)
    float* plane_dc; float* plane_c; int64_t p_frame; int pitch;   // plane_c NULL in the no-confidence path
    int il;                                           // planes interleaved sector by sector in ONE array at plane_dc (SolveArgs::il)
    int skip_c;                                       // the exported map doubles as plane c (first row pass reads it there): plane_c is not written
};

// grid = (ceil(groups/128), H, frames), block 128: a thread takes 4 consecutive view columns, the groups being aligned
// to the ROI origin (X = fx % 4 - 4 + 4 * group) so that the plane stores are aligned float4.  Only ROI pixels need the
// disparity; the confidence export needs every pixel.
__host__ __device__ inline int upscale_groups(int W, int fx) { return (W - ((fx & 3) - 4) + 3) / 4; }

__global__ void __launch_bounds__(128) upscale_prep_kernel(const UpArgs a)
{
    const int grp = blockIdx.x * blockDim.x + threadIdx.x, Y = blockIdx.y, frame = blockIdx.z;
    const int X0 = (a.fx & 3) - 4 + 4 * grp;
    if (X0 >= a.W) return;
    const bool row_in = Y >= a.fy && Y < a.fy + a.fh;
    const bool any_roi = row_in && X0 + 3 >= a.fx && X0 < a.fx + a.fw;
    const bool want_conf = a.conf && a.conf_lo;
    if (!any_roi && !want_conf) return;
    const Lin cy = lin_coeff(Y, a.scale_y, a.dh, false);
    const int y0 = min(max(cy.s, 0), a.dh - 1), y1 = min(max(cy.s + 1, 0), a.dh - 1);
    const float b1 = cy.f, b0 = __fsub_rn(1.f, b1);
    const float* c0 = a.conf_lo ? a.conf_lo + (size_t)frame * a.cl_frame + (size_t)y0 * a.cl_step : nullptr;
    const float* c1 = a.conf_lo ? a.conf_lo + (size_t)frame * a.cl_frame + (size_t)y1 * a.cl_step : nullptr;
    const int16_t* d0 = a.disp + (size_t)frame * a.d_frame + (size_t)y0 * a.d_step;
    const int16_t* d1 = a.disp + (size_t)frame * a.d_frame + (size_t)y1 * a.d_step;
    float cval[4] = { 0.f, 0.f, 0.f, 0.f }, dcv[4] = { 0.f, 0.f, 0.f, 0.f };
    bool roi[4];
    // exactly 2x horizontally (the sample's downscale mode) and away from the image border: the four pixels read source
    // columns base .. base+3 with the fixed coefficients 0.75 / 0.25 that lin_coeff would produce, so every source value
    // is loaded once instead of twice and no coefficient arithmetic runs.  Same float operations in the same order.
    const bool fast2x = (a.W == 2 * a.dw) && X0 >= 1 && X0 + 3 <= a.W - 2;
    if (fast2x) {
        const int odd = X0 & 1, base = (X0 >> 1) - 1 + odd;
        // pixel q takes taps (o, o + 1): even X0 -> (0,1)(1,2)(1,2)(2,3), odd X0 -> (0,1)(0,1)(1,2)(1,2); a1 = 0.75 for even X
        const float a1e = 0.75f, a1o = 0.25f;
        float hc0[4], hc1[4], hd0[4], hd1[4];
        auto taps = [&](auto* r, float (&h)[4]) {
            const float v0 = (float)r[base], v1 = (float)r[base + 1], v2 = (float)r[base + 2], v3 = odd ? 0.f : (float)r[base + 3];
            if (!odd) {
                h[0] = __fadd_rn(__fmul_rn(v0, 1.f - a1e), __fmul_rn(v1, a1e)); h[1] = __fadd_rn(__fmul_rn(v1, 1.f - a1o), __fmul_rn(v2, a1o));
                h[2] = __fadd_rn(__fmul_rn(v1, 1.f - a1e), __fmul_rn(v2, a1e)); h[3] = __fadd_rn(__fmul_rn(v2, 1.f - a1o), __fmul_rn(v3, a1o));
            } else {
                h[0] = __fadd_rn(__fmul_rn(v0, 1.f - a1o), __fmul_rn(v1, a1o)); h[1] = __fadd_rn(__fmul_rn(v0, 1.f - a1e), __fmul_rn(v1, a1e));
                h[2] = __fadd_rn(__fmul_rn(v1, 1.f - a1o), __fmul_rn(v2, a1o)); h[3] = __fadd_rn(__fmul_rn(v1, 1.f - a1e), __fmul_rn(v2, a1e));
            }
        };
        if (a.conf_lo) { taps(c0, hc0); taps(c1, hc1); }
        if (any_roi) { taps(d0, hd0); taps(d1, hd1); }
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int X = X0 + q;
            roi[q] = row_in && X >= a.fx && X < a.fx + a.fw;
            if (a.conf_lo) cval[q] = __fadd_rn(__fmul_rn(hc0[q], b0), __fmul_rn(hc1[q], b1));
            if (roi[q]) {
                const int16_t up = cvt_s16(__fadd_rn(__fmul_rn(hd0[q], b0), __fmul_rn(hd1[q], b1)));
                const float dv = (float)cvt_s16(__fmul_rn((float)up, a.x_ratio));
                dcv[q] = a.plane_c ? __fmul_rn(cval[q], dv) : dv;
            }
        }
    } else
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        const int X = X0 + q;
        roi[q] = row_in && X >= a.fx && X < a.fx + a.fw;
        if (X < 0 || X >= a.W) continue;
        if (!roi[q] && !want_conf) continue;
        const Lin cx = lin_coeff(X, a.scale_x, a.dw, true);
        if (a.conf_lo) cval[q] = __fadd_rn(__fmul_rn(lin_h(c0, cx, a.dw), b0), __fmul_rn(lin_h(c1, cx, a.dw), b1));
        if (roi[q]) {
            const int16_t up = cvt_s16(__fadd_rn(__fmul_rn(lin_h(d0, cx, a.dw), b0), __fmul_rn(lin_h(d1, cx, a.dw), b1)));   // resize -> CV_16S
            const float dv = (float)cvt_s16(__fmul_rn((float)up, a.x_ratio));                                            // disp * x_ratio -> CV_16S (244/273)
            dcv[q] = a.plane_c ? __fmul_rn(cval[q], dv) : dv;
        }
    }
    if (want_conf) {
        float* E = a.conf + (size_t)frame * a.conf_frame + (size_t)Y * a.conf_step + X0;
        if (X0 >= 0 && X0 + 3 < a.W && (((uintptr_t)E) & 15) == 0) *(float4*)E = make_float4(cval[0], cval[1], cval[2], cval[3]);
        else
#pragma unroll
            for (int q = 0; q < 4; ++q) if (X0 + q >= 0 && X0 + q < a.W) E[q] = cval[q];
    }
    if (!any_roi) return;
    // X0 - fx is a non-negative multiple of 4 here: the group lies inside one 8-column sector of the ROI-cropped planes
    const int pc0 = X0 - a.fx;
    const size_t po = a.il ? (size_t)frame * (2 * a.p_frame) + (size_t)(Y - a.fy) * (2 * a.pitch) + (size_t)(pc0 >> 3) * 16 + (pc0 & 7)
                           : (size_t)frame * a.p_frame + (size_t)(Y - a.fy) * a.pitch + pc0;
    float* pc = a.il ? a.plane_dc + 8 : a.plane_c;
    if (roi[0] && roi[3]) {
        if (a.plane_c && !a.skip_c) *(float4*)(pc + po) = make_float4(cval[0], cval[1], cval[2], cval[3]);
        *(float4*)(a.plane_dc + po) = make_float4(dcv[0], dcv[1], dcv[2], dcv[3]);
    } else {
#pragma unroll
        for (int q = 0; q < 4; ++q)
            if (roi[q]) { if (a.plane_c && !a.skip_c) pc[po + q] = cval[q]; a.plane_dc[po + q] = dcv[q]; }
    }
}

// ---- standalone FGS: interleaved source -> float planes, float planes -> interleaved dst ------
// A thread converts 4 consecutive pixels: 4 * CN elements of T = CN * sizeof(T) 32-bit words on the interleaved side (word
// accesses when the row is 4-byte aligned, element accesses otherwise), one float4 per channel on the plane side (plane rows
// are padded to a multiple of 8 floats, so whole float4 are always in bounds).  grid = (ceil(w / 1024), h), block 256
template <typename T> __device__ __forceinline__ T cvt_out(float v);
template <> __device__ __forceinline__ uint8_t cvt_out<uint8_t>(float v) { return cvt_u8(v); }
template <> __device__ __forceinline__ int16_t cvt_out<int16_t>(float v) { return cvt_s16(v); }
template <> __device__ __forceinline__ float cvt_out<float>(float v) { return v; }

template <typename T> __device__ __forceinline__ float word_elem(const uint32_t* w, int e);
template <> __device__ __forceinline__ float word_elem<uint8_t>(const uint32_t* w, int e) { return (float)((w[e >> 2] >> (8 * (e & 3))) & 0xffu); }
template <> __device__ __forceinline__ float word_elem<int16_t>(const uint32_t* w, int e) { return (float)(short)(w[e >> 1] >> (16 * (e & 1))); }
template <> __device__ __forceinline__ float word_elem<float>(const uint32_t* w, int e) { return __uint_as_float(w[e]); }

template <typename T, int CN>
__global__ void __launch_bounds__(256) unpack_kernel(const T* __restrict__ src, int64_t sstep_bytes, int w,
                                                     float* __restrict__ planes, int64_t plane_stride, int pitch)
{
    constexpr int NW = CN * (int)sizeof(T);                          // words per 4 pixels
    const int j0 = 4 * (blockIdx.x * blockDim.x + threadIdx.x), i = blockIdx.y;
    if (j0 >= w) return;
    const T* row = (const T*)((const uint8_t*)src + (size_t)i * sstep_bytes) + (size_t)j0 * CN;
    float v[CN][4];
    if (j0 + 3 < w && (((uintptr_t)row) & 3) == 0) {
        uint32_t wd[NW];
#pragma unroll
        for (int k = 0; k < NW; ++k) wd[k] = __ldg((const uint32_t*)row + k);
#pragma unroll
        for (int q = 0; q < 4; ++q)
#pragma unroll
            for (int ch = 0; ch < CN; ++ch) v[ch][q] = word_elem<T>(wd, q * CN + ch);
    } else {
#pragma unroll
        for (int q = 0; q < 4; ++q)
#pragma unroll
            for (int ch = 0; ch < CN; ++ch) v[ch][q] = (j0 + q < w) ? (float)row[q * CN + ch] : 0.f;
    }
#pragma unroll
    for (int ch = 0; ch < CN; ++ch)
        *(float4*)(planes + (size_t)ch * plane_stride + (size_t)i * pitch + j0) = make_float4(v[ch][0], v[ch][1], v[ch][2], v[ch][3]);
}

template <typename T, int CN>
__global__ void __launch_bounds__(256) pack_kernel(const float* __restrict__ planes, int64_t plane_stride, int pitch, int w,
                                                   T* __restrict__ dst, int64_t dstep_bytes)
{
    constexpr int NW = CN * (int)sizeof(T);
    const int j0 = 4 * (blockIdx.x * blockDim.x + threadIdx.x), i = blockIdx.y;
    if (j0 >= w) return;
    T* row = (T*)((uint8_t*)dst + (size_t)i * dstep_bytes) + (size_t)j0 * CN;
    T o[4 * CN];
#pragma unroll
    for (int ch = 0; ch < CN; ++ch) {
        const float4 f = *(const float4*)(planes + (size_t)ch * plane_stride + (size_t)i * pitch + j0);
        o[0 * CN + ch] = cvt_out<T>(f.x); o[1 * CN + ch] = cvt_out<T>(f.y); o[2 * CN + ch] = cvt_out<T>(f.z); o[3 * CN + ch] = cvt_out<T>(f.w);
    }
    if (j0 + 3 < w && (((uintptr_t)row) & 3) == 0) {
        uint32_t wd[NW];
#pragma unroll
        for (int k = 0; k < NW; ++k) {
            if (sizeof(T) == 1) wd[k] = (uint32_t)(uint8_t)o[4 * k] | ((uint32_t)(uint8_t)o[4 * k + 1] << 8) | ((uint32_t)(uint8_t)o[4 * k + 2] << 16) | ((uint32_t)(uint8_t)o[4 * k + 3] << 24);
            else if (sizeof(T) == 2) wd[k] = (uint32_t)(uint16_t)o[2 * k] | ((uint32_t)(uint16_t)o[2 * k + 1] << 16);
            else wd[k] = __float_as_uint((float)o[k]);
        }
#pragma unroll
        for (int k = 0; k < NW; ++k) ((uint32_t*)row)[k] = wd[k];
    } else {
#pragma unroll
        for (int q = 0; q < 4; ++q)
#pragma unroll
            for (int ch = 0; ch < CN; ++ch) if (j0 + q < w) row[q * CN + ch] = o[q * CN + ch];
    }
}

}  // namespace dwls
