// dwls_passes.cuh — the horizontal and vertical FGS sweeps as single-pass on-chip kernels
// (reference: HorizontalPass_ParBody fgs_filter.cpp:466-476, VerticalPass_ParBody 484-584,
// final normalisation disparity_filters.cpp:295-296).
//
// Algorithmic traffic per sweep and pixel: 4 B weight + 4*S B read + 4*S B write; the Thomas
// intermediates the reference spills to interD (fgs_filter.cpp:75) never leave the SM.
#pragma once
#include <cuda.h>
#include "dwls_solver.cuh"

namespace dwls {

// elements per thread in the row solve: 8 (one 32-byte sector per plane) up to 2048 columns, 16 for wider rows so
// that the reduced system still has at most 256 rows (level 2 in registers, 8 rows per lane)
constexpr int MAXS = 4;      // planes solved together (fgs_filter.cpp:184: channels <= 4)

struct SolveArgs {
    const float* weight;     // Chor (row pass) or Cvert (column pass), ROI-cropped workspace plane
    int64_t w_frame;         // floats between frames
    const float* in[MAXS];   // S planes (frame 0)
    float* out[MAXS];
    int64_t p_frame;         // floats between frames of a plane
    // first row pass only: plane s may be read from somewhere with its own strides (the exported confidence map standing in
    // for plane c); 0 = p_frame / pitch.  Rows there must be 16-byte aligned and hold `pitch` readable floats.
    int64_t in_frame[MAXS]; int in_pitch[MAXS];
    // first row pass of the confidence path (hsolve_kernel DC): plane 0 = d*c is FORMED here as c * float(L) from the left
    // disparity map (at the ROI origin; rows 16-byte aligned, `pitch` readable elements) and plane 1 = c, instead of being read
    const int16_t* dl16; int64_t dl16_frame, dl16_step;     // elements
    int pitch;               // floats between rows, multiple of 8, shared by all workspace planes
    int w, h;
    float lam;
    // il != 0: the S = 2 planes live in ONE array (in[0] / out[0]) interleaved sector by sector, [row][col / 8][plane][8]:
    // a row of both planes is one contiguous 2 * pitch floats (one bulk copy in the row pass) and an 8-column strip of both
    // planes is 64 contiguous bytes per row (one TMA box row instead of two 32-byte ones in the column pass)
    int il;
    // fused epilogue of the last column pass
    int16_t* out16;          // points at the ROI origin inside the full-size output
    int64_t out16_step;      // elements between rows
    int64_t out16_frame;     // elements between frames
};

enum { FIN_FLOAT = 0, FIN_WLS = 1, FIN_S16 = 2 };

// 256-bit global store (sm_100: STG.256), p 32-byte aligned
__device__ __forceinline__ void st_global_v8(float* p, float a, float b, float c, float d, float e, float f, float g, float h)
{
    asm volatile("st.global.v8.f32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d), "f"(e), "f"(f), "f"(g), "f"(h) : "memory");
}

// OpenCV saturate_cast<short>(float): cvRound (ties-to-even; NaN / |v|>=2^31 -> INT_MIN) then clamp
__device__ __forceinline__ int16_t cvt_s16(float v)
{
    short r;
    asm("cvt.rni.sat.s16.f32 %0, %1;" : "=h"(r) : "f"(v));
    return (fabsf(v) < 2147483648.0f) ? (int16_t)r : (int16_t)-32768;
}
__device__ __forceinline__ uint8_t cvt_u8(float v)
{
    if (!(fabsf(v) < 2147483648.0f)) return 0;
    int i = __float2int_rn(v);
    return (uint8_t)max(0, min(255, i));
}
// disparity_filters.cpp:295: d*c~ .mul( 1 / (c~ + EPS) ), EPS = 1e-43f (denormal, line 47)
__device__ __forceinline__ float wls_den(float c) { return __fadd_rn(fmaxf(c, 0.0f), __int_as_float(71)); }   // 71 * 2^-149 == 1e-43f
// fl(1 / den) for 0 < den < 2^64, subnormal den included, without a branch: a tiny den is first scaled by 2^64 (exact), the
// reciprocal taken as MUFU.RCP + one Newton step (the correctly rounded reciprocal up to ties 2^-22 apart, which is what
// the reference's IEEE division returns) and scaled back (exact, or overflow to +inf exactly where 1 / den overflows).
// Zero confidence next to sharp guide edges leaves the filtered confidence at 0 or subnormal over whole regions, so this
// range is common, not an edge case.
__device__ __forceinline__ float wls_rcp_scaled(float den)
{
    const float sc = den < 8.6736174e-19f ? 18446744073709551616.0f : 1.0f;       // 2^-60 : 2^64
    const float ds = __fmul_rn(den, sc);
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(ds));
    r = fmaf(r, fmaf(-ds, r, 1.0f), r);
    return __fmul_rn(r, sc);
}
__device__ __forceinline__ bool wls_den_huge(float den) { return !(den < 18446744073709551616.0f); }   // never with confidences <= 255 (NaN counts as huge)
__device__ __forceinline__ float wls_normalise(float dc, float c)
{
    const float den = wls_den(c);
    return __fmul_rn(dc, wls_den_huge(den) ? __fdiv_rn(1.0f, den) : wls_rcp_scaled(den));
}

// ---------------------------------------------------------------------------------------------
// row pass: persistent CTAs, each loops over image rows (row index runs over all frames of the
// launch).  blockDim.x = K threads (multiple of 32, K*8 >= w); thread t owns columns [8t, 8t+8).
// The weight row and the S plane rows of the NEXT rows are fetched by the TMA engine
// (cp.async.bulk, completion on an mbarrier) into a ring of HST shared-memory stages while the
// current row is being solved from registers, so bytes in flight do not depend on occupancy.
// grid = (min(rows_total, CTAs that fit), 1)
// dynamic smem = HST * (1+S) * pitch floats | 2 * (4+2S) * l2pad_size(K) floats | HST mbarriers
// ---------------------------------------------------------------------------------------------
#ifndef DWLS_PACK16
#define DWLS_PACK16 0      // final column pass: 1 = four columns trade rows and a lane stores 4 CV_16S values with one 64-bit store.
                           // Measured (round 2, 1080p): 7.26 us / frame with it, 7.15 without - the pass is bound by its TMA tile
                           // loads (32-byte box rows), not by its stores - so it stays off; kept for A/B runs.
#endif
#ifndef DWLS_H16MINB
#define DWLS_H16MINB 2     // CTAs per SM the 16-elements-per-thread row pass (rows of 2049..4096 pixels) is compiled for
                           // (3 was measured in round 2: 80 registers with spills, 4K row pass 33.0 -> 36.5 us)
#endif
#ifndef DWLS_HST
#define DWLS_HST 2
#endif
#ifndef DWLS_HMINB
#define DWLS_HMINB 4
#endif
#ifndef DWLS_HL2BUF
#define DWLS_HL2BUF 1
#endif
constexpr int HL2BUF = DWLS_HL2BUF; // level-2 scratch copies (1 needs a third barrier per row, 2 does not)
constexpr int HST = DWLS_HST;      // shared-memory stages per CTA
constexpr int HMINB = DWLS_HMINB;  // CTAs per SM the row pass is compiled for (register cap)

template <int S, int HM>
__host__ __device__ inline size_t hsolve_smem_bytes(int K, int pitch)
{
    return sizeof(float) * ((size_t)HST * (1 + S) * pitch + HL2BUF * (size_t)(4 + 2 * S) * l2pad_size(K) + 32 * (2 + S) * HM) + sizeof(uint64_t) * HST;
}

// DC: see SolveArgs::dl16 - the slot of plane 0 in a stage holds the int16 row of the left disparity map (half of it)
template <int S, int MAXT, int HM, bool IL = false, bool DC = false>
__global__ void __launch_bounds__(MAXT, (MAXT <= 256 ? (HM == 8 ? HMINB : DWLS_H16MINB) : 1)) hsolve_kernel(const SolveArgs a, const int rows_total)
{
    extern __shared__ __align__(128) float sm[];
    const int K = blockDim.x, M2 = K >> 5, KP = l2pad_size(K);
    const int t = threadIdx.x, j0 = t * HM;
    // interleaved planes: threads whose bit 1 is set hold plane 1 in f[0] and plane 0 in f[1] (the planes are solved
    // alike), which spreads a warp's 64-byte-strided reads over all banks instead of half of them
    const int sel = IL ? ((t >> 1) & 1) : 0;
    float* stage0 = sm;                                           // [HST][1+S][pitch]
    float* l2base = sm + (size_t)HST * (1 + S) * a.pitch;         // [2][(4+2S)*KP]
    float* stash = l2base + HL2BUF * (4 + 2 * S) * KP;                 // [(2+S)*HM][32]: warp 0 parks its level-1 state here during level 2
    uint64_t* full = (uint64_t*)(stash + 32 * (2 + S) * HM);      // [HST]
    const uint32_t row_bytes = (uint32_t)a.pitch * 4u;
    const int n_mine = (rows_total - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;

    // my i-th row is number blockIdx.x + i * gridDim.x; its (frame, row in the frame) advance by a constant step
    const int step_f = (int)gridDim.x / a.h, step_r = (int)gridDim.x - step_f * a.h;
    auto issue = [&](int i, int frame, int row) {                 // thread 0 only: fetch my i-th row into stage i % HST
        float* dst = stage0 + (size_t)(i % HST) * (1 + S) * a.pitch;
        uint64_t* bar = full + (i % HST);
        static_assert(!DC || (S == 2 && HM == 8 && !IL), "DC is the first row pass of the confidence path");
        mbar_arrive_expect_tx(bar, DC ? 2 * row_bytes + row_bytes / 2 : (1 + S) * row_bytes);
        bulk_g2s(dst, a.weight + (size_t)frame * a.w_frame + (size_t)row * a.pitch, row_bytes, bar);
        if (DC) {
            bulk_g2s(dst + a.pitch, a.dl16 + (size_t)frame * a.dl16_frame + (size_t)row * a.dl16_step, row_bytes / 2, bar);
            bulk_g2s(dst + 2 * (size_t)a.pitch, a.in[1] + (size_t)frame * (a.in_frame[1] ? a.in_frame[1] : a.p_frame) + (size_t)row * (a.in_pitch[1] ? a.in_pitch[1] : a.pitch), row_bytes, bar);
        } else if (IL) bulk_g2s(dst + a.pitch, a.in[0] + (size_t)frame * (S * a.p_frame) + (size_t)row * (S * a.pitch), S * row_bytes, bar);
        else
#pragma unroll
            for (int s = 0; s < S; ++s)
                bulk_g2s(dst + (size_t)(1 + s) * a.pitch, a.in[s] + (size_t)frame * (a.in_frame[s] ? a.in_frame[s] : a.p_frame) + (size_t)row * (a.in_pitch[s] ? a.in_pitch[s] : a.pitch), row_bytes, bar);
    };

    if (t == 0) {
        for (int i = 0; i < HST; ++i) mbar_init(full + i, 1);
        mbar_fence_init();
    }
    __syncthreads();
    const int tma_thread = K > 32 ? 32 : 0;
    int frame = (int)blockIdx.x / a.h, row = (int)blockIdx.x - frame * a.h;
    int pframe = frame, prow = row;                               // the row HST ahead (the one whose fetch is issued next)
    for (int i = 0; i < HST; ++i) {
        if (t == 0 && i < n_mine) issue(i, pframe, prow);
        pframe += step_f; prow += step_r;
        if (prow >= a.h) { prow -= a.h; ++pframe; }
    }

    for (int i = 0; i < n_mine; ++i) {
        const float* st = stage0 + (size_t)(i % HST) * (1 + S) * a.pitch;
        float* A = l2base + (size_t)(HL2BUF > 1 ? (i & 1) : 0) * (4 + 2 * S) * KP;
        float* B = A + KP; float* Bd = B + KP; float* C = Bd + KP; float* F = C + KP; float* Fd = F + S * KP;
        mbar_wait(full + (i % HST), (uint32_t)((i / HST) & 1));

        float c[HM], v[HM], f[S][HM];
        float aprev = 0.f;
        if (j0 < a.w) {
#pragma unroll
            for (int q = 0; q < HM / 4; ++q) {
                const float4 cc = *(const float4*)(st + j0 + 4 * q);
                c[4 * q] = cc.x; c[4 * q + 1] = cc.y; c[4 * q + 2] = cc.z; c[4 * q + 3] = cc.w;
            }
            if (t > 0) aprev = a.lam * st[j0 - 1];
#pragma unroll
            for (int s = 0; s < S; ++s) {
                if (DC && s == 0) continue;
                const float* P = IL ? st + a.pitch + j0 * S + 8 * (s ^ sel) : st + (size_t)(1 + s) * a.pitch + j0;
#pragma unroll
                for (int q = 0; q < HM / 4; ++q) {
                    const float4 ff = *(const float4*)(P + 4 * q);
                    f[s][4 * q] = ff.x; f[s][4 * q + 1] = ff.y; f[s][4 * q + 2] = ff.z; f[s][4 * q + 3] = ff.w;
                }
            }
            if (DC) {
                // d*c = conf .mul( float(disp) )   (disparity_filters.cpp:289-290): one rounding, like the confidence kernel's own product
                const uint4 L4 = *(const uint4*)((const int16_t*)(st + a.pitch) + j0);
                const uint32_t lw[4] = { L4.x, L4.y, L4.z, L4.w };
#pragma unroll
                for (int k = 0; k < HM; ++k) {
                    const int Lk = (k & 1) ? ((int)lw[k >> 1] >> 16) : (int)(short)(lw[k >> 1] & 0xffffu);
                    f[0][k] = __fmul_rn(f[S - 1][k], (float)Lk);
                }
            }
            if (j0 + HM < a.w) {                                       // chunk strictly inside the row: no masks
#pragma unroll
                for (int k = 0; k < HM; ++k) c[k] = a.lam * c[k];
            } else {
#pragma unroll
                for (int k = 0; k < HM; ++k) {
                    c[k] = (j0 + k < a.w - 1) ? a.lam * c[k] : 0.f;  // last weight of a line is 0 (fgs_filter.cpp:614)
                    if (j0 + k >= a.w) {
#pragma unroll
                        for (int s = 0; s < S; ++s) f[s][k] = 0.f;
                    }
                }
            }
        } else {
#pragma unroll
            for (int k = 0; k < HM; ++k) {
                c[k] = 0.f;
#pragma unroll
                for (int s = 0; s < S; ++s) f[s][k] = 0.f;
            }
        }
        const Level1Out<HM, S> o = level1_reduce<HM, S>(c, f, v, aprev);
        const int p = l2pad(t);
        A[p] = o.A; B[p] = o.B;
#pragma unroll
        for (int s = 0; s < S; ++s) F[(s ^ sel) * KP + p] = o.Fs[s];
        if (t > 0) {
            const int pm = l2pad(t - 1);
            C[pm] = o.Cn; Bd[pm] = o.dB;
#pragma unroll
            for (int s = 0; s < S; ++s) Fd[(s ^ sel) * KP + pm] = o.dF[s];
        }
        if (t == K - 1) {
            C[p] = 0.f; Bd[p] = 0.f;
#pragma unroll
            for (int s = 0; s < S; ++s) Fd[s * KP + p] = 0.f;
        }
        __syncthreads();                       // level-2 rows complete; every thread is done reading this stage
        if (t == tma_thread && i + HST < n_mine) issue(i + HST, pframe, prow);      // not the level-2 warp: its solve is the serial section
        pframe += step_f; prow += step_r;
        if (prow >= a.h) { prow -= a.h; ++pframe; }
        if (t < 32) {
            // the level-2 warp needs its registers for the reduced system: park the level-1 state in shared memory meanwhile
#pragma unroll
            for (int k = 0; k < HM; ++k) {
                stash[(k) * 32 + t] = c[k]; stash[(HM + k) * 32 + t] = v[k];
#pragma unroll
                for (int s = 0; s < S; ++s) stash[((2 + s) * HM + k) * 32 + t] = f[s][k];
            }
            level2_solve_warp<S>(A, B, Bd, C, F, Fd, KP, M2, t);
#pragma unroll
            for (int k = 0; k < HM; ++k) {
                c[k] = stash[(k) * 32 + t]; v[k] = stash[(HM + k) * 32 + t];
#pragma unroll
                for (int s = 0; s < S; ++s) f[s][k] = stash[((2 + s) * HM + k) * 32 + t];
            }
        }
        __syncthreads();
        float Uown[S], Uprev[S];
#pragma unroll
        for (int s = 0; s < S; ++s) { Uown[s] = F[(s ^ sel) * KP + p]; Uprev[s] = (t > 0) ? F[(s ^ sel) * KP + l2pad(t - 1)] : 0.f; }
        if (HL2BUF == 1) __syncthreads();       // the single scratch copy is rewritten by the next row
        level1_fixup<HM, S>(c, f, v, Uprev, Uown);
        if (j0 < a.w) {
#pragma unroll
            for (int s = 0; s < S; ++s) {
                float* P = IL ? a.out[0] + (size_t)frame * (S * a.p_frame) + (size_t)row * (S * a.pitch) + j0 * S + 8 * (s ^ sel)
                              : a.out[s] + (size_t)frame * a.p_frame + (size_t)row * a.pitch + j0;
#pragma unroll
                for (int q = 0; q < HM / 8; ++q)                  // full 32-byte sectors, one 256-bit store each
                    if (q == 0 || j0 + 8 * q < a.pitch) st_global_v8(P + 8 * q, f[s][8 * q], f[s][8 * q + 1], f[s][8 * q + 2], f[s][8 * q + 3], f[s][8 * q + 4], f[s][8 * q + 5], f[s][8 * q + 6], f[s][8 * q + 7]);
            }
        }
        frame += step_f; row += step_r;
        if (row >= a.h) { row -= a.h; ++frame; }
    }
}

// ---------------------------------------------------------------------------------------------
// column pass: persistent CTAs (one per SM), each loops over strips of VC adjacent columns (8; 4 or 2 for tall images)
// (strip index runs over all frames of the launch).  Thread (x = tid % 8, k = tid / 8) owns rows
// [k*M, k*M+M) of column x of the strip.  A strip (h rows x 32 bytes, for the weight plane and
// the S data planes) is fetched as 2-D TMA tiles into ONE shared-memory stage; every thread pulls
// its M rows into registers (bank-conflict free because M is odd), after which the stage is
// released and the TMA engine fetches the next strip while this one is solved and stored.
// Every global access is therefore full 32-byte sectors, and neighbouring strips (the other
// sectors of the same 128-byte lines) are in flight on other SMs at the same time.
// blockDim.x = 8 * K, K multiple of 32, K*M >= h.   grid = min(#strips, #SMs)
// dynamic smem = (1+S) * rows_staged * 8 floats | 8 * (4+2S) * l2pad_size(K) floats | mbarrier
// ---------------------------------------------------------------------------------------------
struct VTma {
    int rb;            // rows per TMA box (<= 256)
    int nb;            // boxes per strip: nb * rb >= h
    int strips_x;      // strips per frame = ceil(w / VC)
    int strips_total;  // strips_x * frames
};

// wbuf: staged copies of the weight strip (1, or 2 when it is double-buffered)
template <int M, int S, int VC>
__host__ __device__ inline size_t vsolve_smem_bytes(int K, int rows_staged, int wbuf)
{
    return sizeof(float) * ((size_t)(wbuf + S) * rows_staged * VC + (size_t)VC * (4 + 2 * S) * l2pad_size(K) + (size_t)VC * 32 * (2 + S) * M) + 32;
}

// WB: the weight plane is strip-major ([strip][row][VC], written so by weights_kernel) and a strip arrives with one bulk
// copy into one of two buffers, requested a whole strip ahead (the weights are only read by level 1, so their buffer is
// free again long before the data planes' stage is); otherwise it is row-major like the data planes and arrives as TMA
// boxes together with them.
template <int M, int S, int VC, int FIN, int MAXT, bool WB, bool IL = false>
__global__ void __launch_bounds__(MAXT) vsolve_kernel(const SolveArgs a, const VTma g, const __grid_constant__ CUtensorMap tm_w,
                                                      const __grid_constant__ CUtensorMap tm_p0, const __grid_constant__ CUtensorMap tm_p1)
{
    extern __shared__ __align__(128) float sm[];
    const int K = blockDim.x / VC, M2 = K >> 5, KP = l2pad_size(K);
    const int SYS = (4 + 2 * S) * KP;
    const int rows_staged = g.nb * g.rb;
    constexpr int NW = WB ? 2 : 1;
    float* wstage = sm;                                           // [NW][rows_staged][VC]
    float* stage = sm + (size_t)NW * rows_staged * VC;            // [S][rows_staged][VC]
    float* l2base = stage + (size_t)S * rows_staged * VC;         // [VC][SYS]
    float* stash = l2base + (size_t)VC * SYS;                      // [(2+S)*M][32*VC]: level-2 warps park their level-1 state
    uint64_t* full = (uint64_t*)(stash + (size_t)VC * 32 * (2 + S) * M);
    uint64_t* wfull = full + 1;                                   // [2]
    uint64_t* udone = full + 3;                                   // every thread arrives once it has read its level-2 results
    const int t = threadIdx.x, x = t % VC, k = t / VC, r0 = k * M;
    // interleaved planes: every other pair of chunks holds plane 1 in f[0] and plane 0 in f[1], so that the four chunks of a
    // warp read four different bank groups of the [row][plane][VC] stage
    const int sel = IL ? ((k >> 1) & 1) : 0;
    float* A = l2base + x * SYS; float* B = A + KP; float* Bd = B + KP; float* C = Bd + KP;
    float* F = C + KP; float* Fd = F + S * KP;
    const int n_mine = (g.strips_total - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
    const int tma_thread = ((int)blockDim.x > 32 * VC) ? 32 * VC : 0;
    const uint32_t tx_bytes = (uint32_t)(((WB ? 0 : rows_staged) + S * rows_staged) * VC * sizeof(float));

    // my i-th strip is number blockIdx.x + i * gridDim.x; its (frame, strip in the frame) advance by a constant step
    const int step_f = (int)gridDim.x / g.strips_x, step_s = (int)gridDim.x - step_f * g.strips_x;
    auto issue = [&](int frame, int strip) {                      // thread 0: fetch the data planes of a strip
        const int col0 = strip * VC;
        mbar_arrive_expect_tx(full, tx_bytes);
        for (int b = 0; b < g.nb; ++b) {
            const int grow = frame * a.h + b * g.rb;
            if (!WB) tma_load_2d(wstage + (size_t)b * g.rb * VC, &tm_w, col0, grow, full);
            if (IL) { tma_load_2d(stage + (size_t)(b * g.rb) * (VC * S), &tm_p0, col0 * S, grow, full); continue; }   // [row][plane][VC]
            tma_load_2d(stage + (size_t)(b * g.rb) * VC, &tm_p0, col0, grow, full);
            if (S > 1) tma_load_2d(stage + (size_t)(rows_staged + b * g.rb) * VC, &tm_p1, col0, grow, full);
        }
    };
    auto issue_w = [&](int i, int frame, int strip) {             // thread 0: fetch the weight strip of my i-th strip
        const uint32_t bytes = (uint32_t)(a.h * VC * sizeof(float));
        mbar_arrive_expect_tx(wfull + (i & 1), bytes);
        bulk_g2s(wstage + (size_t)(i & 1) * rows_staged * VC, a.weight + (size_t)frame * a.w_frame + (size_t)strip * ((size_t)a.h * VC), bytes, wfull + (i & 1));
    };

    if (t == 0) { mbar_init(full, 1); mbar_init(wfull, 1); mbar_init(wfull + 1, 1); mbar_init(udone, blockDim.x); mbar_fence_init(); }
    __syncthreads();
    int frame = (int)blockIdx.x / g.strips_x, strip = (int)blockIdx.x - frame * g.strips_x;
    if (t == 0 && n_mine > 0) { if (WB) issue_w(0, frame, strip); issue(frame, strip); }

    for (int i = 0; i < n_mine; ++i) {
        const int col = strip * VC + x;
        const bool live = col < a.w;
        int nframe = frame + step_f, nstrip = strip + step_s;    // my next strip
        if (nstrip >= g.strips_x) { nstrip -= g.strips_x; ++nframe; }
        if (WB) {
            if (t == 0 && i + 1 < n_mine) issue_w(i + 1, nframe, nstrip);   // its buffer was last read by level 1 of strip i-1
            mbar_wait(wfull + (i & 1), (uint32_t)((i >> 1) & 1));
        }
        mbar_wait(full, (uint32_t)(i & 1));
        float c[M], v[M], f[S][M];
        float aprev = 0.f;
        {
            const float* sw = wstage + (size_t)(WB ? (i & 1) : 0) * rows_staged * VC + (size_t)r0 * VC + x;
            const float* sp = stage + (size_t)r0 * (IL ? VC * S : VC) + x;
            constexpr int SPM = IL ? VC * S : VC;                        // floats between rows of a staged plane
            const int sps = IL ? VC : rows_staged * VC;                  // floats between planes
            if (live && r0 + M < a.h) {                                  // chunk strictly inside the column: no masks
#pragma unroll
                for (int m = 0; m < M; ++m) {
                    c[m] = a.lam * sw[m * VC];
#pragma unroll
                    for (int s = 0; s < S; ++s) f[s][m] = sp[(s ^ sel) * sps + m * SPM];
                }
            } else {
#pragma unroll
                for (int m = 0; m < M; ++m) {
                    const bool in = live && (r0 + m < a.h);
                    c[m] = (live && (r0 + m < a.h - 1)) ? a.lam * sw[m * VC] : 0.f;              // last row weight 0 (fgs_filter.cpp:658-660)
#pragma unroll
                    for (int s = 0; s < S; ++s) f[s][m] = in ? sp[(s ^ sel) * sps + m * SPM] : 0.f;
                }
            }
            if (live && k > 0 && r0 - 1 < a.h - 1) aprev = a.lam * sw[-VC];
        }
        const Level1Out<M, S> o = level1_reduce<M, S>(c, f, v, aprev);
        const int p = l2pad(k);
        if (i > 0) mbar_wait(udone, (uint32_t)((i - 1) & 1));
        A[p] = o.A; B[p] = o.B;
#pragma unroll
        for (int s = 0; s < S; ++s) F[(s ^ sel) * KP + p] = o.Fs[s];
        if (k > 0) {
            const int pm = l2pad(k - 1);
            C[pm] = o.Cn; Bd[pm] = o.dB;
#pragma unroll
            for (int s = 0; s < S; ++s) Fd[(s ^ sel) * KP + pm] = o.dF[s];
        }
        if (k == K - 1) {
            C[p] = 0.f; Bd[p] = 0.f;
#pragma unroll
            for (int s = 0; s < S; ++s) Fd[s * KP + p] = 0.f;
        }
        __syncthreads();                       // level-2 rows complete; the stage has been read by everyone
        // the next strip's tiles are requested by the first thread that is NOT in a level-2 warp: the dozen TMA instructions
        // would otherwise sit at the head of system 0's solve, the serial section every other warp is waiting for
        if (t == tma_thread && i + 1 < n_mine) issue(nframe, nstrip);
        {
            const int warp = t >> 5, nwarps = blockDim.x >> 5, lane = t & 31;
            if (warp < VC) {
                // the level-2 warps need their registers for the reduced systems: park the level-1 state meanwhile
                const int ST = 32 * VC;
#pragma unroll
                for (int m = 0; m < M; ++m) {
                    stash[m * ST + t] = c[m]; stash[(M + m) * ST + t] = v[m];
#pragma unroll
                    for (int s = 0; s < S; ++s) stash[((2 + s) * M + m) * ST + t] = f[s][m];
                }
                for (int sys = warp; sys < VC; sys += nwarps) {
                    float* A2 = l2base + sys * SYS;
                    level2_solve_warp<S>(A2, A2 + KP, A2 + 2 * KP, A2 + 3 * KP, A2 + 4 * KP, A2 + (4 + S) * KP, KP, M2, lane);
                }
#pragma unroll
                for (int m = 0; m < M; ++m) {
                    c[m] = stash[m * ST + t]; v[m] = stash[(M + m) * ST + t];
#pragma unroll
                    for (int s = 0; s < S; ++s) f[s][m] = stash[((2 + s) * M + m) * ST + t];
                }
            }
        }
        __syncthreads();
        float Uown[S], Uprev[S];
#pragma unroll
        for (int s = 0; s < S; ++s) { Uown[s] = F[(s ^ sel) * KP + p]; Uprev[s] = (k > 0) ? F[(s ^ sel) * KP + l2pad(k - 1)] : 0.f; }
        mbar_arrive(udone);                    // split barrier: the level-2 scratch may be overwritten once everybody has arrived,
                                               // which the next strip checks just before it writes there (nobody waits here)
        level1_fixup<M, S>(c, f, v, Uprev, Uown);
        constexpr bool PAIR = (FIN == FIN_FLOAT && S == 2 && !IL && VC % 2 == 0 && M % 2 == 1);
        if (PAIR) {
            // neighbouring columns swap every other row, so that a lane stores TWO adjacent columns of one row with a 64-bit
            // store: even columns keep rows 0, 2, .., odd columns rows 1, 3, ..; the last row (M - 1) pairs plane 0 (even
            // columns) with plane 1 (odd columns).  Half as many store instructions for the LSU queue (9 instead of 18).
            const bool odd = x & 1;
            const int nrow = min(M, a.h - r0);
            float* Ob = a.out[0] + (size_t)frame * a.p_frame + (size_t)r0 * a.pitch + (col - (x & 1));
            const size_t dplane = (size_t)(a.out[1] - a.out[0]);
#pragma unroll
            for (int s = 0; s < S; ++s) {
#pragma unroll
                for (int m = 0; m + 1 < M; m += 2) {
                    const float got = __shfl_xor_sync(0xffffffffu, odd ? f[s][m] : f[s][m + 1], 1);
                    const float lo = odd ? got : f[s][m], hi = odd ? f[s][m + 1] : got;
                    const int mr = m + (odd ? 1 : 0);
                    if (mr < nrow) *(float2*)(Ob + s * dplane + (size_t)mr * a.pitch) = make_float2(lo, hi);
                }
            }
            const float got = __shfl_xor_sync(0xffffffffu, odd ? f[0][M - 1] : f[1][M - 1], 1);
            const float lo = odd ? got : f[0][M - 1], hi = odd ? f[1][M - 1] : got;
            if (M - 1 < nrow) *(float2*)(Ob + (odd ? dplane : 0) + (size_t)(M - 1) * a.pitch) = make_float2(lo, hi);
        }
        if (FIN == FIN_FLOAT) {
            if (live && !PAIR) {
                const size_t off0 = IL ? (size_t)frame * (S * a.p_frame) + (size_t)r0 * (S * a.pitch) + (size_t)(col - x) * S + x
                                       : (size_t)frame * a.p_frame + (size_t)r0 * a.pitch + col;
                const int opitch = IL ? S * a.pitch : a.pitch;
                const int nrow = min(M, a.h - r0);                       // rows of this chunk inside the image (<= 0 for padding chunks)
#pragma unroll
                for (int s = 0; s < S; ++s) {
                    float* O = IL ? a.out[0] + off0 + (s ^ sel) * VC : a.out[s] + off0;
                    if (nrow == M) {
#pragma unroll
                        for (int m = 0; m < M; ++m) O[m * opitch] = f[s][m];
                    } else {
#pragma unroll
                        for (int m = 0; m < M; ++m) if (m < nrow) O[m * opitch] = f[s][m];
                    }
                }
            }
        } else {
            // fused epilogue of the last column pass: normalise (confidence path), round to CV_16S, store into the ROI of the
            // full-size output.  Every thread runs the arithmetic and the shuffles (padding chunks and columns past the ROI
            // hold zeros); only the stores are predicated.
            const int nrow = min(M, a.h - r0);
            if (IL && sel) {                                             // this chunk holds the planes the other way round
#pragma unroll
                for (int m = 0; m < M; ++m) { const float q = f[0][m]; f[0][m] = f[S - 1][m]; f[S - 1][m] = q; }
            }
            if (FIN == FIN_WLS) {
                // one range test for the chunk, then straight-line code (the exact division is for denominators no
                // confidence map produces)
                bool huge = false;
#pragma unroll
                for (int m = 0; m < M; ++m) { f[S - 1][m] = wls_den(f[S - 1][m]); huge |= wls_den_huge(f[S - 1][m]); }
                if (!huge) {
#pragma unroll
                    for (int m = 0; m < M; ++m) f[0][m] = __fmul_rn(f[0][m], wls_rcp_scaled(f[S - 1][m]));
                } else {
#pragma unroll
                    for (int m = 0; m < M; ++m) f[0][m] = __fmul_rn(f[0][m], __fdiv_rn(1.0f, f[S - 1][m]));
                }
            }
            int16_t* O = a.out16 + (size_t)frame * a.out16_frame + (size_t)r0 * a.out16_step + col;
            // 8-column strips with an 8-byte aligned output: four neighbouring columns trade rows so that a lane stores FOUR
            // adjacent CV_16S values of one row with one 64-bit store (two packed exchanges per 4 rows instead of 4 narrow
            // stores per lane: the LSU queue, not the arithmetic, is what this pass waits for)
            const bool packed = DWLS_PACK16 && VC == 8 && (((uintptr_t)a.out16 | (uintptr_t)(a.out16_step * 2) | (uintptr_t)(a.out16_frame * 2)) & 7) == 0;   // uniform
            if (packed) {
                const int col4 = col - (x & 3);
                const bool grp_ok = col4 + 3 < a.w;                      // the same for the four lanes of a group
                const bool odd = x & 1, hi = x & 2;
#pragma unroll
                for (int g4 = 0; g4 + 3 < M; g4 += 4) {
                    const uint32_t r0v = (uint16_t)cvt_s16(f[0][g4]), r1v = (uint16_t)cvt_s16(f[0][g4 + 1]), r2v = (uint16_t)cvt_s16(f[0][g4 + 2]),
                                   r3v = (uint16_t)cvt_s16(f[0][g4 + 3]);
                    // stage 1 (lane ^ 1): even lanes collect row g4 (and g4+2) of their column pair, odd lanes row g4+1 (g4+3)
                    const uint32_t snd01 = odd ? r0v : r1v, snd23 = odd ? r2v : r3v;
                    const uint32_t got01 = __shfl_xor_sync(0xffffffffu, snd01, 1), got23 = __shfl_xor_sync(0xffffffffu, snd23, 1);
                    const uint32_t w01 = odd ? (got01 | (r1v << 16)) : (r0v | (got01 << 16));
                    const uint32_t w23 = odd ? (got23 | (r3v << 16)) : (r2v | (got23 << 16));
                    // stage 2 (lane ^ 2): the low pair of lanes keeps rows g4 / g4+1, the high pair rows g4+2 / g4+3
                    const uint32_t got = __shfl_xor_sync(0xffffffffu, hi ? w01 : w23, 2);
                    const uint32_t keep = hi ? w23 : w01;
                    const int row = g4 + (odd ? 1 : 0) + (hi ? 2 : 0);
                    if (grp_ok) {
                        if (row < nrow) *(uint2*)(O - (x & 3) + (size_t)row * a.out16_step) = hi ? make_uint2(got, keep) : make_uint2(keep, got);
                    } else if (live) {
                        if (g4 < nrow) O[(size_t)g4 * a.out16_step] = (int16_t)r0v;
                        if (g4 + 1 < nrow) O[(size_t)(g4 + 1) * a.out16_step] = (int16_t)r1v;
                        if (g4 + 2 < nrow) O[(size_t)(g4 + 2) * a.out16_step] = (int16_t)r2v;
                        if (g4 + 3 < nrow) O[(size_t)(g4 + 3) * a.out16_step] = (int16_t)r3v;
                    }
                }
#pragma unroll
                for (int m = M - (M & 3); m < M; ++m)
                    if (live && m < nrow) O[(size_t)m * a.out16_step] = cvt_s16(f[0][m]);
            } else if (live) {
                if (nrow == M) {
#pragma unroll
                    for (int m = 0; m < M; ++m) O[(size_t)m * a.out16_step] = cvt_s16(f[0][m]);
                } else {
#pragma unroll
                    for (int m = 0; m < M; ++m)
                        if (m < nrow) O[(size_t)m * a.out16_step] = cvt_s16(f[0][m]);
                }
            }
        }
        frame = nframe; strip = nstrip;
    }
}

// ---------------------------------------------------------------------------------------------
// tall images (more than 2176 rows): a column no longer fits the shared-memory stage of vsolve_kernel, so the
// column pass runs as a ROW pass over the transposed planes (hsolve_kernel handles lines of up to 8192 unknowns):
// transpose -> hsolve -> transpose back, the latter with the same fused epilogues as the last column pass.
// 32 x 32 tiles through shared memory, both sides coalesced.  in: [rows][in_pitch], out: [cols][out_pitch].
// grid = (ceil(cols/32), ceil(rows/32), frames), block = 256
// ---------------------------------------------------------------------------------------------
struct TrArgs {
    const float* in[2]; int64_t in_frame; int in_pitch;
    float* out[2]; int64_t out_frame; int out_pitch;
    int rows, cols;                                              // of the input
    int16_t* out16; int64_t out16_step, out16_frame;             // FIN != FIN_FLOAT: [cols][rows] at the ROI origin
};

template <int S, int FIN>
__global__ void __launch_bounds__(256) transpose_kernel(const TrArgs a)
{
    __shared__ float tile[S][32][33];
    const int frame = blockIdx.z, c0 = blockIdx.x * 32, r0 = blockIdx.y * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
#pragma unroll
    for (int s = 0; s < S; ++s) {
        const float* I = a.in[s] + (size_t)frame * a.in_frame;
#pragma unroll
        for (int k = ty; k < 32; k += 8) {
            const int r = r0 + k, c = c0 + tx;
            tile[s][k][tx] = (r < a.rows && c < a.cols) ? I[(size_t)r * a.in_pitch + c] : 0.f;
        }
    }
    __syncthreads();
#pragma unroll
    for (int k = ty; k < 32; k += 8) {
        const int orow = c0 + k, ocol = r0 + tx;                 // output coordinates
        if (orow >= a.cols || ocol >= a.rows) continue;
        if (FIN == FIN_FLOAT) {
#pragma unroll
            for (int s = 0; s < S; ++s) a.out[s][(size_t)frame * a.out_frame + (size_t)orow * a.out_pitch + ocol] = tile[s][tx][k];
        } else {
            const float v = (FIN == FIN_WLS) ? wls_normalise(tile[0][tx][k], tile[S - 1][tx][k]) : tile[0][tx][k];
            a.out16[(size_t)frame * a.out16_frame + (size_t)orow * a.out16_step + ocol] = cvt_s16(v);
        }
    }
}

}  // namespace dwls
