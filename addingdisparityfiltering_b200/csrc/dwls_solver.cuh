// dwls_solver.cuh — tridiagonal line solver shared by the row (horizontal) and column (vertical)
// passes of the Fast Global Smoother (reference: modules/ximgproc/src/fgs_filter.cpp:439-464 row
// solve, 484-584 column solve).
//
// The reference runs Thomas' algorithm serially along each line.  A line of 1080..3840 unknowns
// is far too long a dependency chain for a GPU thread, so each line is solved by substructuring
// (Schur complement) in two levels, all on-chip:
//   level 1  every thread owns M consecutive unknowns in REGISTERS, eliminates its M-1 interior
//            unknowns (local Thomas with two spike right-hand sides) and emits one row of a
//            reduced tridiagonal system that couples only the per-thread separators;
//   level 2  one warp per reduced system repeats the same step on chunks held in shared memory
//            and is left with a 32-unknown tridiagonal system, one row per lane, which it solves
//            by warp-shuffle parallel cyclic reduction (5 steps);
//   fix-up   separators are broadcast back through shared memory and every thread finishes
//            its interior unknowns with two FMAs per plane.
// All S planes of a call (d*c and c for the WLS filter, the channels of an FGS source) share
// the weights, the factorisation and the spikes.
//
// System per line (a_j = lam*C[j-1], c_j = lam*C[j], b_j = 1 - a_j - c_j, C <= 0 so b_j >= 1):
//      a_j u[j-1] + b_j u[j] + c_j u[j+1] = f[j]
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace dwls {

// reciprocal of a float known to be a normal number >= 1: MUFU.RCP + one Newton step.
// (.ftz keeps it a single MUFU: without it ptxas wraps the instruction in denormal range checks;
// neither the argument nor the result can be denormal here.)
__device__ __forceinline__ float rcp_nr(float x)
{
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return fmaf(r, fmaf(-x, r, 1.0f), r);
}

// pivots of both levels use the bare MUFU.RCP: relative error <= 2^-23, one float rounding's worth next to the dozens
// of roundings a solve already accumulates.  Measured at 1080p against the CPU reference restatement: 99.9924 % of the CV_16S outputs
// identical, 100 % within 1 LSB (with a Newton step per pivot: 99.9930 %), for two dependent FMAs less per step.
__device__ __forceinline__ float rcp_l2(float x) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }

// ---- mbarrier / bulk-copy (TMA engine) primitives ------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_LOOP:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra WAIT_DONE;\n"
        "bra WAIT_LOOP;\n"
        "WAIT_DONE:\n"
        "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
// 1-D bulk copy global -> shared, completion counted in bytes on an mbarrier (16-byte aligned, size % 16 == 0)
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)), "l"(src),
                 "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}

// 2-D tiled TMA load (tensor map in kernel parameter space), completion counted in bytes on an mbarrier
__device__ __forceinline__ void tma_load_2d(void* dst, const void* tmap, int c0, int c1, uint64_t* bar)
{
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(
                     smem_u32(dst)),
                 "l"(tmap), "r"(c0), "r"(c1), "r"(smem_u32(bar))
                 : "memory");
}

// padded index for level-2 shared arrays (breaks power-of-two lane strides)
__device__ __forceinline__ int l2pad(int i) { return i + (i >> 5); }
__host__ __device__ inline int l2pad_size(int n) { return n + (n >> 5) + 1; }

// ---------------------------------------------------------------------------------------------
// level 1: registers.  On entry c[i] = lam*C of element i, f[s][i] = right-hand sides,
// aprev = lam*C of the element before the chunk (0 for the first chunk of a line).
// On exit, for the interior i = 0..M-2:  u_i = f[s][i] - v[i]*U_prev - c[i]*U_own
// (f holds x, c holds w) and the return values are this chunk's separator row pieces.
// ---------------------------------------------------------------------------------------------
template <int M, int S>
struct Level1Out {
    float A, B, Fs[S];        // own contributions to reduced row k
    float dB, Cn, dF[S];      // contributions of this chunk to reduced row k-1 (the previous chunk's)
};

template <int M, int S>
__device__ __forceinline__ Level1Out<M, S> level1_reduce(float (&c)[M], float (&f)[S][M], float (&v)[M], float aprev)
{
    static_assert(M >= 2, "chunk needs an interior");
    Level1Out<M, S> o;
    // separator coefficients before c[] is overwritten
    const float a_s = c[M - 2], c_s = c[M - 1];
    const float b_s = 1.0f - a_s - c_s;
    // forward elimination over the interior
    float a = aprev, ct_prev = 0.f, vt_prev = 0.f, ft_prev[S];
#pragma unroll
    for (int s = 0; s < S; ++s) ft_prev[s] = 0.f;
#pragma unroll
    for (int i = 0; i < M - 1; ++i) {
        const float ci = c[i];
        const float den = (1.0f - a - ci) - a * ct_prev;
        const float r = rcp_l2(den);
        const float ct = ci * r;
        const float vt = (i == 0) ? a * r : -(a * vt_prev) * r;
#pragma unroll
        for (int s = 0; s < S; ++s) { const float ft = (f[s][i] - a * ft_prev[s]) * r; f[s][i] = ft; ft_prev[s] = ft; }
        c[i] = ct; v[i] = vt;
        ct_prev = ct; vt_prev = vt;
        a = ci;
    }
    // backward substitution: x (in f), v, w (in c)
    float w_next = c[M - 2], v_next = v[M - 2];      // w_{M-2} = ct_{M-2}
#pragma unroll
    for (int i = M - 3; i >= 0; --i) {
        const float ct = c[i];
#pragma unroll
        for (int s = 0; s < S; ++s) f[s][i] = fmaf(-ct, f[s][i + 1], f[s][i]);
        v_next = fmaf(-ct, v_next, v[i]); v[i] = v_next;
        w_next = -ct * w_next; c[i] = w_next;
    }
    // separator row k:  A U_{k-1} + B U_k + C U_{k+1} = F ; own part uses the last interior element
    o.A = -a_s * v[M - 2];
    o.B = b_s - a_s * c[M - 2];
#pragma unroll
    for (int s = 0; s < S; ++s) o.Fs[s] = f[s][M - 1] - a_s * f[s][M - 2];
    // what this chunk's first interior element contributes to the PREVIOUS separator row (its c_s = aprev)
    o.dB = -aprev * v[0];
    o.Cn = -aprev * c[0];
#pragma unroll
    for (int s = 0; s < S; ++s) o.dF[s] = -aprev * f[s][0];
    return o;
}

template <int M, int S>
__device__ __forceinline__ void level1_fixup(const float (&c)[M], float (&f)[S][M], const float (&v)[M],
                                             const float (&Uprev)[S], const float (&Uown)[S])
{
#pragma unroll
    for (int i = 0; i < M - 1; ++i)
#pragma unroll
        for (int s = 0; s < S; ++s) f[s][i] = fmaf(-c[i], Uown[s], fmaf(-v[i], Uprev[s], f[s][i]));
#pragma unroll
    for (int s = 0; s < S; ++s) f[s][M - 1] = Uown[s];
}

// ---------------------------------------------------------------------------------------------
// level 2, generic trip count (wide rows only, M2 > 8): chunks stay in shared memory
// (32*M2 rows at padded index l2pad(i)).  Row k is assembled from two writers: chunk k stores
// A, B, F and chunk k+1 stores C, Bd, Fd (its first element's share), so B+Bd and F+Fd are the
// actual diagonal and right-hand side.  Result: the solution overwrites F[s].
// ---------------------------------------------------------------------------------------------
template <int S>
__device__ __noinline__ void level2_solve_loop(float* A, const float* B, const float* Bd, float* C, float* F, const float* Fd,
                                  int fstride, int M2, int lane)
{
    const int s0 = lane * M2, q = s0 + M2 - 1;          // interior s0..q-1, separator q
    float xL[S], vL = -1.f, wL = 0.f;                    // last interior element (defaults: empty interior)
    float xF[S], vF = 0.f, wF = -1.f;                    // first interior element
#pragma unroll
    for (int s = 0; s < S; ++s) { xL[s] = 0.f; xF[s] = 0.f; }
    if (M2 > 1) {
        float ct_prev = 0.f, vt_prev = 0.f, ft_prev[S];
#pragma unroll
        for (int s = 0; s < S; ++s) ft_prev[s] = 0.f;
        for (int i = s0; i < q; ++i) {
            const int p = l2pad(i);
            const float a = A[p];
            const float r = 1.0f / ((B[p] + Bd[p]) - a * ct_prev);
            const float ct = C[p] * r;
            const float vt = (i == s0) ? a * r : -(a * vt_prev) * r;
#pragma unroll
            for (int s = 0; s < S; ++s) { const float ft = ((F[s * fstride + p] + Fd[s * fstride + p]) - a * ft_prev[s]) * r; F[s * fstride + p] = ft; ft_prev[s] = ft; }
            C[p] = ct; A[p] = vt; ct_prev = ct; vt_prev = vt;
        }
        float w_next = ct_prev, v_next = vt_prev, x_next[S];
#pragma unroll
        for (int s = 0; s < S; ++s) x_next[s] = ft_prev[s];
        wL = w_next; vL = v_next;
#pragma unroll
        for (int s = 0; s < S; ++s) xL[s] = x_next[s];
        for (int i = q - 2; i >= s0; --i) {
            const int p = l2pad(i);
            const float ct = C[p];
#pragma unroll
            for (int s = 0; s < S; ++s) { x_next[s] = fmaf(-ct, x_next[s], F[s * fstride + p]); F[s * fstride + p] = x_next[s]; }
            v_next = fmaf(-ct, v_next, A[p]); A[p] = v_next;
            w_next = -ct * w_next; C[p] = w_next;
        }
        vF = v_next; wF = w_next;
#pragma unroll
        for (int s = 0; s < S; ++s) xF[s] = x_next[s];
    }
    // this lane's row of the 32-row system; the next lane's first element enters through C_q
    const int pq = l2pad(q);
    const float Aq = A[pq], Bq = B[pq] + Bd[pq], Cq = C[pq];
    const unsigned FULL = 0xffffffffu;
    float nvF = __shfl_down_sync(FULL, vF, 1), nwF = __shfl_down_sync(FULL, wF, 1);
    if (lane == 31) { nvF = 0.f; nwF = 0.f; }
    float a = -Aq * vL;
    float b = Bq - Aq * wL - Cq * nvF;
    float c = -Cq * nwF;
    float g[S];
#pragma unroll
    for (int s = 0; s < S; ++s) {
        float nxF = __shfl_down_sync(FULL, xF[s], 1);
        if (lane == 31) nxF = 0.f;
        g[s] = (F[s * fstride + pq] + Fd[s * fstride + pq]) - Aq * xL[s] - Cq * nxF;
    }
    // parallel cyclic reduction across the 32 lanes
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        float am = __shfl_up_sync(FULL, a, d), bm = __shfl_up_sync(FULL, b, d), cm = __shfl_up_sync(FULL, c, d);
        float ap = __shfl_down_sync(FULL, a, d), bp = __shfl_down_sync(FULL, b, d), cp = __shfl_down_sync(FULL, c, d);
        const bool lo = lane < d, hi = lane + d > 31;
        if (lo) { am = 0.f; bm = 1.f; cm = 0.f; }
        if (hi) { ap = 0.f; bp = 1.f; cp = 0.f; }
        const float k1 = a / bm, k2 = c / bp;
#pragma unroll
        for (int s = 0; s < S; ++s) {
            float gm = __shfl_up_sync(FULL, g[s], d), gp = __shfl_down_sync(FULL, g[s], d);
            if (lo) gm = 0.f;
            if (hi) gp = 0.f;
            g[s] = g[s] - gm * k1 - gp * k2;
        }
        b = b - cm * k1 - ap * k2;
        a = -am * k1;
        c = -cp * k2;
    }
    float U[S], Up[S];
#pragma unroll
    for (int s = 0; s < S; ++s) {
        U[s] = g[s] / b;
        Up[s] = __shfl_up_sync(FULL, U[s], 1);
        if (lane == 0) Up[s] = 0.f;
    }
    for (int i = s0; i < q; ++i) {
        const int p = l2pad(i);
        const float vv = A[p], ww = C[p];
#pragma unroll
        for (int s = 0; s < S; ++s) F[s * fstride + p] = fmaf(-ww, U[s], fmaf(-vv, Up[s], F[s * fstride + p]));
    }
#pragma unroll
    for (int s = 0; s < S; ++s) F[s * fstride + pq] = U[s];
}

// ---------------------------------------------------------------------------------------------
// level 2: one warp solves one reduced tridiagonal system of 32*M2 rows held in shared memory at
// padded index l2pad(i).  Row k is assembled from two writers: chunk k stores A, B, F and chunk
// k+1 stores C, Bd, Fd (its first element's share), so B+Bd and F+Fd are the actual diagonal and
// right-hand side.  Each lane pulls its M2 consecutive rows into registers, eliminates its M2-1
// interior rows (same step as level 1), the 32 separator rows are solved by parallel cyclic
// reduction over warp shuffles, and the solution overwrites F[s].
// This is the only serial section of a sweep (the other warps of the CTA wait at a barrier), so
// it is written for latency: all loads issued up front, compile-time trip counts, reciprocals
// by MUFU + one Newton step (every pivot is >= 1: the Schur complements stay diagonally dominant).
// ---------------------------------------------------------------------------------------------
template <int S, int M2>
__device__ __forceinline__ void level2_solve_regs(float* A, const float* B, const float* Bd, float* C, float* F, const float* Fd,
                                                  int fstride, int lane)
{
    const unsigned FULL = 0xffffffffu;
    const int s0 = lane * M2;
    float a[M2], b[M2], c[M2], f[S][M2];
#pragma unroll
    for (int i = 0; i < M2; ++i) {
        const int p = l2pad(s0 + i);
        a[i] = A[p]; b[i] = B[p] + Bd[p]; c[i] = C[p];
#pragma unroll
        for (int s = 0; s < S; ++s) f[s][i] = F[s * fstride + p] + Fd[s * fstride + p];
    }
    // interior rows 0..M2-2: forward elimination with the left spike in a[], then back substitution
    // leaving u_i = f[s][i] - a[i]*U_prev - c[i]*U_own
    float xL[S], vL = -1.f, wL = 0.f, xF[S], vF = 0.f, wF = -1.f;     // defaults describe an empty interior (M2 == 1)
#pragma unroll
    for (int s = 0; s < S; ++s) { xL[s] = 0.f; xF[s] = 0.f; }
    if (M2 > 1) {
#pragma unroll
        for (int i = 0; i < M2 - 1; ++i) {
            const float ai = a[i];
            const float den = (i == 0) ? b[0] : fmaf(-ai, c[i - 1], b[i]);
            const float r = rcp_l2(den);
            c[i] = c[i] * r;
            a[i] = (i == 0) ? ai * r : -(ai * a[i - 1]) * r;
#pragma unroll
            for (int s = 0; s < S; ++s) f[s][i] = ((i == 0) ? f[s][0] : fmaf(-ai, f[s][i - 1], f[s][i])) * r;
        }
        vL = a[M2 - 2]; wL = c[M2 - 2];
#pragma unroll
        for (int s = 0; s < S; ++s) xL[s] = f[s][M2 - 2];
#pragma unroll
        for (int i = M2 - 3; i >= 0; --i) {
            const float ct = c[i];
#pragma unroll
            for (int s = 0; s < S; ++s) f[s][i] = fmaf(-ct, f[s][i + 1], f[s][i]);
            a[i] = fmaf(-ct, a[i + 1], a[i]);
            c[i] = -ct * c[i + 1];
        }
        vF = a[0]; wF = c[0];
#pragma unroll
        for (int s = 0; s < S; ++s) xF[s] = f[s][0];
    }
    // this lane's row of the 32-row system; the next lane's first interior row enters through cq
    const float aq = a[M2 - 1], bq = b[M2 - 1], cq = c[M2 - 1];
    float nvF = __shfl_down_sync(FULL, vF, 1), nwF = __shfl_down_sync(FULL, wF, 1);
    if (lane == 31) { nvF = 0.f; nwF = 0.f; }
    float pa = -aq * vL;
    float pb = bq - aq * wL - cq * nvF;
    float pc = -cq * nwF;
    float g[S];
#pragma unroll
    for (int s = 0; s < S; ++s) {
        float nxF = __shfl_down_sync(FULL, xF[s], 1);
        if (lane == 31) nxF = 0.f;
        g[s] = f[s][M2 - 1] - aq * xL[s] - cq * nxF;
    }
    // parallel cyclic reduction across the 32 lanes (rows beyond either end act as identity rows)
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        float am = __shfl_up_sync(FULL, pa, d), bm = __shfl_up_sync(FULL, pb, d), cm = __shfl_up_sync(FULL, pc, d);
        float ap = __shfl_down_sync(FULL, pa, d), bp = __shfl_down_sync(FULL, pb, d), cp = __shfl_down_sync(FULL, pc, d);
        const bool lo = lane < d, hi = lane + d > 31;
        if (lo) { am = 0.f; bm = 1.f; cm = 0.f; }
        if (hi) { ap = 0.f; bp = 1.f; cp = 0.f; }
        const float k1 = pa * rcp_l2(bm), k2 = pc * rcp_l2(bp);
#pragma unroll
        for (int s = 0; s < S; ++s) {
            float gm = __shfl_up_sync(FULL, g[s], d), gp = __shfl_down_sync(FULL, g[s], d);
            if (lo) gm = 0.f;
            if (hi) gp = 0.f;
            g[s] = fmaf(-gp, k2, fmaf(-gm, k1, g[s]));
        }
        pb = fmaf(-ap, k2, fmaf(-cm, k1, pb));
        pa = -am * k1;
        pc = -cp * k2;
    }
    const float rb = rcp_l2(pb);
#pragma unroll
    for (int s = 0; s < S; ++s) {
        const float U = g[s] * rb;
        float Up = __shfl_up_sync(FULL, U, 1);
        if (lane == 0) Up = 0.f;
#pragma unroll
        for (int i = 0; i < M2 - 1; ++i) F[s * fstride + l2pad(s0 + i)] = fmaf(-c[i], U, fmaf(-a[i], Up, f[s][i]));
        F[s * fstride + l2pad(s0 + M2 - 1)] = U;
    }
}

template <int S>
__device__ __forceinline__ void level2_solve_warp(float* A, const float* B, const float* Bd, float* C, float* F, const float* Fd,
                                                  int fstride, int M2, int lane)
{
    switch (M2) {
    case 1: level2_solve_regs<S, 1>(A, B, Bd, C, F, Fd, fstride, lane); break;
    case 2: level2_solve_regs<S, 2>(A, B, Bd, C, F, Fd, fstride, lane); break;
    case 3: level2_solve_regs<S, 3>(A, B, Bd, C, F, Fd, fstride, lane); break;
    case 4: level2_solve_regs<S, 4>(A, B, Bd, C, F, Fd, fstride, lane); break;
    case 5: level2_solve_regs<S, 5>(A, B, Bd, C, F, Fd, fstride, lane); break;
    case 6: level2_solve_regs<S, 6>(A, B, Bd, C, F, Fd, fstride, lane); break;
    case 7: level2_solve_regs<S, 7>(A, B, Bd, C, F, Fd, fstride, lane); break;
    case 8: level2_solve_regs<S, 8>(A, B, Bd, C, F, Fd, fstride, lane); break;
    default: level2_solve_loop<S>(A, B, Bd, C, F, Fd, fstride, M2, lane); break;
    }
}

}  // namespace dwls
