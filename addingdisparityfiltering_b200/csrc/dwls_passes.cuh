// dwls_passes.cuh — the horizontal and vertical FGS sweeps as single-pass on-chip kernels
// (reference: HorizontalPass_ParBody fgs_filter.cpp:466-476, VerticalPass_ParBody 484-584,
// final normalisation disparity_filters.cpp:295-296).
//
// Algorithmic traffic per sweep and pixel: 4 B weight + 4*S B read + 4*S B write; the Thomas
// intermediates the reference spills to interD (fgs_filter.cpp:75) never leave the SM.
#pragma once
#include "dwls_solver.cuh"

namespace dwls {

constexpr int HM = 8;        // elements per thread in the row solve (two float4)
constexpr int MAXS = 4;      // planes solved together (fgs_filter.cpp:184: channels <= 4)

struct SolveArgs {
    const float* weight;     // Chor (row pass) or Cvert (column pass), ROI-cropped workspace plane
    int64_t w_frame;         // floats between frames
    const float* in[MAXS];   // S planes (frame 0)
    float* out[MAXS];
    int64_t p_frame;         // floats between frames of a plane
    int pitch;               // floats between rows, multiple of 8, shared by all workspace planes
    int w, h;
    float lam;
    // fused epilogue of the last column pass
    int16_t* out16;          // points at the ROI origin inside the full-size output
    int64_t out16_step;      // elements between rows
    int64_t out16_frame;     // elements between frames
};

enum { FIN_FLOAT = 0, FIN_WLS = 1, FIN_S16 = 2 };

// OpenCV saturate_cast<short>(float): cvRound (ties-to-even; NaN / |v|>=2^31 -> INT_MIN) then clamp
__device__ __forceinline__ int16_t cvt_s16(float v)
{
    if (!(fabsf(v) < 2147483648.0f)) return (int16_t)-32768;
    int i = __float2int_rn(v);
    return (int16_t)max(-32768, min(32767, i));
}
__device__ __forceinline__ uint8_t cvt_u8(float v)
{
    if (!(fabsf(v) < 2147483648.0f)) return 0;
    int i = __float2int_rn(v);
    return (uint8_t)max(0, min(255, i));
}
// disparity_filters.cpp:295: d*c~ .mul( 1 / (c~ + EPS) ), EPS = 1e-43f (denormal, line 47)
__device__ __forceinline__ float wls_normalise(float dc, float c)
{
    const float den = __fadd_rn(fmaxf(c, 0.0f), __int_as_float(71));   // 71 * 2^-149 == 1e-43f
    const float rec = __fdiv_rn(1.0f, den);
    return __fmul_rn(dc, rec);
}

// ---------------------------------------------------------------------------------------------
// row pass: one CTA per image row, blockDim.x = K threads (multiple of 32, K*8 >= w)
// grid = (h, frames)
// ---------------------------------------------------------------------------------------------
template <int S, int MAXT>
__global__ void __launch_bounds__(MAXT) hsolve_kernel(const SolveArgs a)
{
    extern __shared__ float sm[];
    const int K = blockDim.x, M2 = K >> 5, KP = l2pad_size(K);
    float* A = sm; float* B = A + KP; float* Bd = B + KP; float* C = Bd + KP;
    float* F = C + KP; float* Fd = F + S * KP;
    const int t = threadIdx.x, row = blockIdx.x, frame = blockIdx.y;
    const int j0 = t * HM;
    const size_t roff = (size_t)row * a.pitch + j0;
    const float* W = a.weight + (size_t)frame * a.w_frame + roff;

    float c[HM], v[HM], f[S][HM];
    float aprev = 0.f;
    if (j0 < a.w) {
        const float4 c0 = __ldg((const float4*)W), c1 = __ldg((const float4*)W + 1);
        c[0] = c0.x; c[1] = c0.y; c[2] = c0.z; c[3] = c0.w; c[4] = c1.x; c[5] = c1.y; c[6] = c1.z; c[7] = c1.w;
        if (t > 0) aprev = a.lam * __ldg(W - 1);
#pragma unroll
        for (int s = 0; s < S; ++s) {
            const float* P = a.in[s] + (size_t)frame * a.p_frame + roff;
            const float4 f0 = *(const float4*)P, f1 = *((const float4*)P + 1);
            f[s][0] = f0.x; f[s][1] = f0.y; f[s][2] = f0.z; f[s][3] = f0.w; f[s][4] = f1.x; f[s][5] = f1.y; f[s][6] = f1.z; f[s][7] = f1.w;
        }
#pragma unroll
        for (int i = 0; i < HM; ++i) {
            c[i] = (j0 + i < a.w - 1) ? a.lam * c[i] : 0.f;      // last weight of a line is 0 (fgs_filter.cpp:614)
            if (j0 + i >= a.w) {
#pragma unroll
                for (int s = 0; s < S; ++s) f[s][i] = 0.f;
            }
        }
    } else {
#pragma unroll
        for (int i = 0; i < HM; ++i) {
            c[i] = 0.f;
#pragma unroll
            for (int s = 0; s < S; ++s) f[s][i] = 0.f;
        }
    }
    const Level1Out<HM, S> o = level1_reduce<HM, S>(c, f, v, aprev);
    const int p = l2pad(t);
    A[p] = o.A; B[p] = o.B;
#pragma unroll
    for (int s = 0; s < S; ++s) F[s * KP + p] = o.Fs[s];
    if (t > 0) {
        const int pm = l2pad(t - 1);
        C[pm] = o.Cn; Bd[pm] = o.dB;
#pragma unroll
        for (int s = 0; s < S; ++s) Fd[s * KP + pm] = o.dF[s];
    }
    if (t == K - 1) {
        C[p] = 0.f; Bd[p] = 0.f;
#pragma unroll
        for (int s = 0; s < S; ++s) Fd[s * KP + p] = 0.f;
    }
    __syncthreads();
    if (t < 32) level2_solve_warp<S>(A, B, Bd, C, F, Fd, KP, M2, t);
    __syncthreads();
    float Uown[S], Uprev[S];
#pragma unroll
    for (int s = 0; s < S; ++s) { Uown[s] = F[s * KP + p]; Uprev[s] = (t > 0) ? F[s * KP + l2pad(t - 1)] : 0.f; }
    level1_fixup<HM, S>(c, f, v, Uprev, Uown);
    if (j0 < a.w) {
#pragma unroll
        for (int s = 0; s < S; ++s) {
            float* P = a.out[s] + (size_t)frame * a.p_frame + roff;
            *(float4*)P = make_float4(f[s][0], f[s][1], f[s][2], f[s][3]);
            *((float4*)P + 1) = make_float4(f[s][4], f[s][5], f[s][6], f[s][7]);
        }
    }
}

// ---------------------------------------------------------------------------------------------
// column pass: one CTA per strip of COLS adjacent columns; thread (x = tid % COLS, k = tid / COLS)
// owns rows [k*M, k*M+M) of column x0+x.  blockDim.x = COLS * K, K multiple of 32, K*M >= h.
// grid = (ceil(w / COLS), frames)
// ---------------------------------------------------------------------------------------------
template <int M, int S, int COLS, int FIN, int MAXT>
__global__ void __launch_bounds__(MAXT) vsolve_kernel(const SolveArgs a)
{
    extern __shared__ float sm[];
    const int K = blockDim.x / COLS, M2 = K >> 5, KP = l2pad_size(K);
    const int SYS = (4 + 2 * S) * KP;                       // floats per reduced system
    const int t = threadIdx.x, x = t % COLS, k = t / COLS, frame = blockIdx.y;
    const int col = blockIdx.x * COLS + x, r0 = k * M;
    float* A = sm + x * SYS; float* B = A + KP; float* Bd = B + KP; float* C = Bd + KP;
    float* F = C + KP; float* Fd = F + S * KP;
    const bool live = col < a.w;
    const size_t off0 = (size_t)r0 * a.pitch + col;
    const float* W = a.weight + (size_t)frame * a.w_frame + off0;

    float c[M], v[M], f[S][M];
    float aprev = 0.f;
#pragma unroll
    for (int i = 0; i < M; ++i) {
        const bool in = live && (r0 + i < a.h);
        c[i] = (live && (r0 + i < a.h - 1)) ? a.lam * __ldg(W + (size_t)i * a.pitch) : 0.f;   // last row weight 0 (658-660)
#pragma unroll
        for (int s = 0; s < S; ++s) f[s][i] = in ? a.in[s][(size_t)frame * a.p_frame + off0 + (size_t)i * a.pitch] : 0.f;
    }
    if (live && k > 0 && r0 - 1 < a.h - 1) aprev = a.lam * __ldg(W - a.pitch);

    const Level1Out<M, S> o = level1_reduce<M, S>(c, f, v, aprev);
    const int p = l2pad(k);
    A[p] = o.A; B[p] = o.B;
#pragma unroll
    for (int s = 0; s < S; ++s) F[s * KP + p] = o.Fs[s];
    if (k > 0) {
        const int pm = l2pad(k - 1);
        C[pm] = o.Cn; Bd[pm] = o.dB;
#pragma unroll
        for (int s = 0; s < S; ++s) Fd[s * KP + pm] = o.dF[s];
    }
    if (k == K - 1) {
        C[p] = 0.f; Bd[p] = 0.f;
#pragma unroll
        for (int s = 0; s < S; ++s) Fd[s * KP + p] = 0.f;
    }
    __syncthreads();
    {
        const int warp = t >> 5, nwarps = blockDim.x >> 5, lane = t & 31;
        for (int sys = warp; sys < COLS; sys += nwarps) {
            float* A2 = sm + sys * SYS;
            level2_solve_warp<S>(A2, A2 + KP, A2 + 2 * KP, A2 + 3 * KP, A2 + 4 * KP, A2 + (4 + S) * KP, KP, M2, lane);
        }
    }
    __syncthreads();
    float Uown[S], Uprev[S];
#pragma unroll
    for (int s = 0; s < S; ++s) { Uown[s] = F[s * KP + p]; Uprev[s] = (k > 0) ? F[s * KP + l2pad(k - 1)] : 0.f; }
    level1_fixup<M, S>(c, f, v, Uprev, Uown);
    if (!live) return;
    if (FIN == FIN_FLOAT) {
#pragma unroll
        for (int i = 0; i < M; ++i)
            if (r0 + i < a.h) {
#pragma unroll
                for (int s = 0; s < S; ++s) a.out[s][(size_t)frame * a.p_frame + off0 + (size_t)i * a.pitch] = f[s][i];
            }
    } else {
        int16_t* O = a.out16 + (size_t)frame * a.out16_frame + (size_t)r0 * a.out16_step + col;
#pragma unroll
        for (int i = 0; i < M; ++i)
            if (r0 + i < a.h) {
                const float val = (FIN == FIN_WLS) ? wls_normalise(f[0][i], f[S - 1][i]) : f[0][i];
                O[(size_t)i * a.out16_step] = cvt_s16(val);
            }
    }
}

}  // namespace dwls
