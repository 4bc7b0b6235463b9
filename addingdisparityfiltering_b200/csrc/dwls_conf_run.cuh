// dwls_conf_run.cuh — confidence stage for small discontinuity radii (R <= 7), the common case
// (createDisparityWLSFilter derives R = ceil(0.33 * wsize) / ceil(0.5 * wsize),
// disparity_filters.cpp:402,408).  Same arithmetic as dwls_conf.cuh (which remains the path for
// larger radii): exact integer window sums, then the reference's rounding sequence
// (boxFilterOp / sqrBoxFilterOp 105-115, ComputeDepthDisc_ParBody 349-373, LRC 306-341).
//
// Structure: a CTA stages the haloed int16 tile of the ROI crop in shared memory once (reflect-101
// border of the CROP resolved while staging), after which no thread talks to another: thread
// (run g, band b) marches down CR_RB rows of 8 adjacent output columns, keeping the vertical
// window sums of its 8 + 2R columns in registers (sum: int32, sum of squares: int64), and forms
// the 8 horizontal window sums of every row by a sliding window over those registers.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <cuda.h>
#include "dwls_conf.cuh"
#include "dwls_passes.cuh"

namespace dwls {

// 3-D tiled TMA load {x, y, frame} (tensor map in kernel parameter space), completion counted in bytes on an mbarrier
__device__ __forceinline__ void tma_load_3d(void* dst, const void* tmap, int c0, int c1, int c2, uint64_t* bar)
{
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];" ::"r"(
                     smem_u32(dst)),
                 "l"(tmap), "r"(c0), "r"(c1), "r"(c2), "r"(smem_u32(bar))
                 : "memory");
}

constexpr int CR_NB = 4;                   // row bands per CTA
constexpr int CR_RB = 16;                  // rows per band
constexpr int CR_TH = CR_NB * CR_RB;       // output rows per CTA
constexpr int CR_T = 128;                  // threads per CTA (>= runs * CR_NB)
constexpr int CR_MAXR = 7;                  // static shared memory (tile + gather buffers) stays under 48 KB

constexpr int CR_HL = 8;                   // tile column c holds crop column tx0 - CR_HL + c
constexpr int CR_GMAX = 29;                // runs of 8 px per tile: 8*G + 2*CR_HL + 7 (TMA realignment slack) <= 256

// dynamic shared memory of conf_run_kernel: tile, per warp 2 x (256 right confidences | 256 words holding right disparities), row table
__host__ __device__ inline size_t conf_run_smem(int r, bool left) { return (size_t)(CR_TH + 2 * r + 1) * 512 + (left ? CR_NB * 2 * 512 * 4 : 0) + (CR_TH + 2 * r + 1) * 4 + 16; }

template <int R>
struct CrGeom {
    static constexpr int RA = (R + 1) & ~1;                 // halo rounded up to even: whole 32-bit words
    static constexpr int PITCH = 256;                       // int16 per shared row
    static constexpr int TR = CR_TH + 2 * R + 1;            // tile rows: y0-R-1 .. y0+TH+R-1
    static constexpr int NW = 4 + RA;                       // 32-bit words a thread reads per row (8 + 2*RA int16)
    static constexpr int NCOL = 8 + 2 * R;                  // columns whose vertical sums a thread keeps
};

__device__ __forceinline__ int cr_reflect(int p, int len)
{
    if (p < 0) p = -p;
    if (p >= len) p = 2 * (len - 1) - p;
    if (p < 0 || p >= len) p = reflect101(p, len);          // crops narrower than the halo
    return p;
}

// element e (compile-time) of a packed int16 word array, sign-extended
template <int NWORDS>
__device__ __forceinline__ int cr_elem(const uint32_t (&w)[NWORDS], int e)
{
    const uint32_t x = w[e >> 1];
    return (e & 1) ? ((int)x >> 16) : (int)(short)(x & 0xffffu);
}

// p = the run's 8 centre columns in the tile (16-byte aligned); w receives columns -RA .. 8+RA-1 around it
template <int NWORDS>
__device__ __forceinline__ void cr_load_row(uint32_t (&w)[NWORDS], const uint16_t* p)
{
    constexpr int H = (NWORDS - 4) / 2;                     // halo words on each side: 0..4
    const uint4 c = *(const uint4*)p;
    w[H] = c.x; w[H + 1] = c.y; w[H + 2] = c.z; w[H + 3] = c.w;
    if (H == 1) { w[0] = *(const uint32_t*)(p - 2); w[5] = *(const uint32_t*)(p + 8); }
    if (H == 2) { const uint2 l = *(const uint2*)(p - 4), r = *(const uint2*)(p + 8); w[0] = l.x; w[1] = l.y; w[6] = r.x; w[7] = r.y; }
    if (H == 3) {
        w[0] = *(const uint32_t*)(p - 6); const uint2 l = *(const uint2*)(p - 4); w[1] = l.x; w[2] = l.y;
        const uint2 r = *(const uint2*)(p + 8); w[7] = r.x; w[8] = r.y; w[9] = *(const uint32_t*)(p + 12);
    }
    if (H == 4) {
        const uint4 l = *(const uint4*)(p - 8), r = *(const uint4*)(p + 8);
        w[0] = l.x; w[1] = l.y; w[2] = l.z; w[3] = l.w; w[8] = r.x; w[9] = r.y; w[10] = r.z; w[11] = r.w;
    }
}

// grid = (ceil(rw / (8*G)), ceil(rh / CR_TH), frames), block = CR_T, G = runs per tile (<= CR_GMAX)
// TMA = true : the tile is fetched by one cp.async.bulk.tensor box {256 x TR} of the int16 disparity map
//              (needs a 16-byte aligned base and row/frame strides that are multiples of 16 bytes), after
//              which only the halo cells that fall outside the crop are rewritten with their reflection;
// TMA = false: any alignment, every thread copies one column pair of the tile.
template <int R, bool LEFT, bool TMA>
__global__ void __launch_bounds__(CR_T, 4) conf_run_kernel(const ConfArgs a, const int G, const __grid_constant__ CUtensorMap tm)
{
    using Gm = CrGeom<R>;
    constexpr int RA = Gm::RA, PITCH = Gm::PITCH, TR = Gm::TR, NW = Gm::NW, NCOL = Gm::NCOL, OFF = RA - R;
    extern __shared__ __align__(128) unsigned char cr_smem[];   // tile | gather buffers (left view) | row table (plain-load path)
    uint16_t* tile = (uint16_t*)cr_smem;
    float* capbuf = (float*)(cr_smem + (size_t)TR * PITCH * 2);
    int* rowy = (int*)(capbuf + (LEFT ? CR_NB * 2 * 512 : 0));
    const int tid = threadIdx.x, frame = blockIdx.z;
    const int tx0 = blockIdx.x * 8 * G, ty0 = blockIdx.y * CR_TH;
    const int64_t sstep = LEFT ? a.dl_step : a.dr_step;
    const int16_t* src = (LEFT ? a.dl + (size_t)frame * a.dl_frame : a.dr + (size_t)frame * a.dr_frame) + (size_t)a.ry * sstep + (LEFT ? a.rx : a.rrx);

    // ---- stage the tile: crop rows ty0-R-1 .., crop columns tx0-CR_HL .., reflect-101 of the crop (168-174) ----
    const int xlo = tx0 - CR_HL, ylo = ty0 - R - 1;                                   // crop coordinates of tile cell (0, 0)
    if (TMA) {
        __shared__ __align__(8) uint64_t bar;
        // the box must start on a 16-byte boundary of the row: fetch from the aligned column below and shift afterwards
        const int gx = (LEFT ? a.rx : a.rrx) + xlo, shift = gx & 7;
        if (tid == 0) { mbar_init(&bar, 1); mbar_fence_init(); }
        __syncthreads();
        if (tid == 0) {
            mbar_arrive_expect_tx(&bar, (uint32_t)(TR * PITCH * 2));
            tma_load_3d(tile, &tm, gx - shift, a.ry + ylo, frame, &bar);
        }
        mbar_wait(&bar, 0);
        if (shift) {
            // in place, a batch of rows at a time: thread t rebuilds 32-bit word t of every row from words t+sw, t+sw+1
            constexpr int NBATCH = 24;
            const int sw = shift >> 1, sh = (shift & 1) * 16;
            const int j0 = min(tid + sw, PITCH / 2 - 1), j1 = min(tid + sw + 1, PITCH / 2 - 1);
            uint32_t* t32 = (uint32_t*)tile;
            for (int r0 = 0; r0 < TR; r0 += NBATCH) {
                uint32_t v[NBATCH];
#pragma unroll
                for (int k = 0; k < NBATCH; ++k)
                    if (r0 + k < TR) v[k] = __funnelshift_r(t32[(r0 + k) * (PITCH / 2) + j0], t32[(r0 + k) * (PITCH / 2) + j1], sh);
                __syncthreads();
#pragma unroll
                for (int k = 0; k < NBATCH; ++k)
                    if (r0 + k < TR) t32[(r0 + k) * (PITCH / 2) + tid] = v[k];
            }
            __syncthreads();
        }
        // cells of the tile outside the crop hold image data or zeros: overwrite the ones the windows reach
        const bool fix_l = xlo < 0, fix_r = xlo + PITCH > a.rw, fix_t = ylo < 0, fix_b = ylo + TR > a.rh;
        if (fix_l || fix_r) {
            // columns -R..-1 and rw..rw+R-1, every tile row
            const int nl = fix_l ? R : 0, nr = fix_r ? min(R, xlo + PITCH - a.rw) : 0, nfix = nl + nr;
            for (int idx = tid; idx < TR * nfix; idx += CR_T) {
                const int ty = idx / nfix, k = idx - ty * nfix;
                const int x = k < nl ? k - R : a.rw + (k - nl);
                const int xs = cr_reflect(x, a.rw);
                if (xs - xlo >= 0 && xs - xlo < PITCH) tile[ty * PITCH + (x - xlo)] = tile[ty * PITCH + (xs - xlo)];
            }
            __syncthreads();
        }
        if (fix_t || fix_b) {
            // rows < 0 and rows rh..rh+R-1, whole rows (after the column fix, so corners come out right)
            const int nt = fix_t ? -ylo : 0, nb = fix_b ? min(R, ylo + TR - a.rh) : 0;
            for (int idx = tid; idx < (nt + nb) * (PITCH / 2); idx += CR_T) {
                const int k = idx / (PITCH / 2), wq = idx - k * (PITCH / 2);
                const int y = k < nt ? ylo + k : a.rh + (k - nt);
                const int ys = cr_reflect(y, a.rh);
                if (ys - ylo >= 0 && ys - ylo < TR) ((uint32_t*)tile)[(y - ylo) * (PITCH / 2) + wq] = ((const uint32_t*)tile)[(ys - ylo) * (PITCH / 2) + wq];
            }
            __syncthreads();
        }
    } else {
        for (int ty = tid; ty < TR; ty += CR_T) rowy[ty] = cr_reflect(ylo + ty, a.rh);
        __syncthreads();
        const int npairs = (8 * G + 2 * CR_HL) >> 1;
        const bool pairs_ok = (((uintptr_t)src & 3) == 0) && ((sstep & 1) == 0);       // xlo is even
        if (tid < npairs) {
            const int xc = xlo + 2 * tid;
            uint32_t* dst = (uint32_t*)tile + tid;
            if (pairs_ok && xc >= 0 && xc + 1 < a.rw) {
                const int16_t* p = src + xc;
#pragma unroll 23
                for (int ty = 0; ty < TR; ++ty) dst[ty * (PITCH / 2)] = __ldg((const uint32_t*)(p + (size_t)rowy[ty] * sstep));
            } else {
                const int16_t* p0 = src + cr_reflect(xc, a.rw);
                const int16_t* p1 = src + cr_reflect(xc + 1, a.rw);
#pragma unroll 8
                for (int ty = 0; ty < TR; ++ty) {
                    const size_t off = (size_t)rowy[ty] * sstep;
                    dst[ty * (PITCH / 2)] = (uint32_t)(uint16_t)__ldg(p0 + off) | ((uint32_t)(uint16_t)__ldg(p1 + off) << 16);
                }
            }
        }
        __syncthreads();
    }

    // ---- per-thread march: warp = band of CR_RB rows, lane = run of 8 columns ----
    const int band = tid >> 5, g = tid & 31;
    const int cx0 = tx0 + 8 * g;                               // first output column of the run (crop coordinates)
    const int yb = ty0 + band * CR_RB;
    if (yb >= a.rh) return;                                    // warp-uniform
    const bool active = g < G && cx0 < a.rw;                   // lanes without a run still serve the LRC gathers of their warp
    if (!LEFT && !active) return;
    const int yend = min(yb + CR_RB, a.rh);
    const int nvalid = min(8, a.rw - cx0);

    int s[NCOL]; long long q[NCOL];
#pragma unroll
    for (int c = 0; c < NCOL; ++c) { s[c] = 0; q[c] = 0; }
    // tile row of crop row y is y - ty0 + R + 1; warm-up: rows yb-R-1 .. yb+R-1 (the first of them is the row the first step removes)
    const uint16_t* tp = tile + (size_t)(band * CR_RB) * PITCH + CR_HL + 8 * (active ? g : 0);   // tile row (yb - R - 1), the run's 8 centre columns
#pragma unroll 1
    for (int k = 0; k < 2 * R + 1; ++k) {
        uint32_t w[NW];
        cr_load_row<NW>(w, tp + (size_t)k * PITCH);
#pragma unroll
        for (int c = 0; c < NCOL; ++c) { const int v = cr_elem<NW>(w, c + OFF); s[c] += v; q[c] += v * v; }
    }

    // left view: the data-dependent gathers of the LRC check (306-341) are issued TRANSPOSED (lane l takes pixels
    // l, l+32, .. of the warp's row segment, so a warp request touches neighbouring addresses) as 4-byte cp.async
    // copies into the warp's shared buffer, one row ahead, and land while the sums of the current row are formed.
    // Misses are issued with src-size 0 (zero fill, no global access) so the loop has no branches.
    auto issue_gather = [&](int y, int buf) {
        // 64-bit row bases once, 32-bit offsets per element (the offsets are bounded by the row length).  All eight left
        // disparities are read before the first copy is issued: every shared-memory load that precedes a cp.async costs
        // three filler instructions (LDS -> LDGSTS hazard), so the copies go out as one uninterrupted run.
        const int16_t* crow = (const int16_t*)tile + (size_t)(y - ylo) * PITCH + CR_HL + g;
        const uint64_t dbase = (uint64_t)(uintptr_t)(a.dr + (size_t)frame * a.dr_frame + (size_t)(a.ry + y) * a.dr_step + a.rrx);
        const uint64_t dal = dbase & ~(uint64_t)3;                       // the copies fetch the aligned word holding the element
        const uint32_t dpar = (uint32_t)(dbase >> 1) & 1u;
        const uint64_t cbase = (uint64_t)(uintptr_t)(a.conf_r + (size_t)frame * a.cr_frame + (size_t)y * a.cr_pitch);
        const int xr = a.rx + tx0 + g - a.rrx;
        const uint32_t cb = smem_u32(capbuf + (band * 2 + buf) * 512 + g);
        int L[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) L[k] = crow[32 * k];
        uint32_t roc[8], sz[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const bool in = 32 * k + g < 8 * G;
            const int ro = xr + 32 * k - (L[k] >> 4);                                              // 331, relative to the right ROI
            const bool ok = in && (unsigned)ro < (unsigned)a.rw;                                   // 332
            roc[k] = ok ? (uint32_t)ro : 0u;
            sz[k] = ok ? 4u : 0u;
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const uint64_t ca = cbase + (uint64_t)(roc[k] * 4u);
            const uint64_t da = dal + (uint64_t)(((roc[k] + dpar) >> 1) * 4u);
            asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(cb + 4 * (32 * k)), "l"(ca), "r"(sz[k]) : "memory");
            asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(cb + 4 * (256 + 32 * k)), "l"(da), "r"(sz[k]) : "memory");
        }
    };
    if (LEFT) { issue_gather(yb, 0); asm volatile("cp.async.commit_group;" ::: "memory"); }

    for (int y = yb; y < yend; ++y, tp += PITCH) {
        const int gx0 = a.rx + cx0, gy = a.ry + y;
        if (LEFT) {
            if (y + 1 < yend) issue_gather(y + 1, (y + 1 - yb) & 1);
            asm volatile("cp.async.commit_group;" ::: "memory");
        }
        uint32_t win[NW], wout[NW];
        cr_load_row<NW>(wout, tp);                               // row y-R-1 leaves
        cr_load_row<NW>(win, tp + (size_t)(2 * R + 1) * PITCH);  // row y+R enters
#pragma unroll
        for (int c = 0; c < NCOL; ++c) {
            const int vi = cr_elem<NW>(win, c + OFF), vo = cr_elem<NW>(wout, c + OFF);
            const int d = vi - vo;
            s[c] += d;
            q[c] += (long long)d * (long long)(vi + vo);         // vi^2 - vo^2
        }
        // horizontal sliding window over the register columns
        int S = 0; long long Q = 0;
#pragma unroll
        for (int c = 0; c <= 2 * R; ++c) { S += s[c]; Q += q[c]; }
        float cv[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const float mean = __fmul_rn((float)S, a.scale_f);
            // Q < 2^52: exact conversion by the 2^52 bias
            const double Qd = __dadd_rn(__longlong_as_double(Q | 0x4330000000000000LL), -4503599627370496.0);
            const float meansq = (float)__dmul_rn(Qd, a.scale_d);
            const float variance = __fsub_rn(meansq, __fmul_rn(mean, mean));
            cv[i] = fmaxf(__fsub_rn(1.0f, __fmul_rn(a.roll_off, variance)), 0.0f);
            if (i < 7) { S += s[i + 2 * R + 1] - s[i]; Q += q[i + 2 * R + 1] - q[i]; }
        }
        if (!LEFT) {
            float* O = a.conf_r + (size_t)frame * a.cr_frame + (size_t)y * a.cr_pitch + cx0;
            if (nvalid == 8) st_global_v8(O, cv[0], cv[1], cv[2], cv[3], cv[4], cv[5], cv[6], cv[7]);
            else
#pragma unroll
                for (int i = 0; i < 8; ++i) if (i < nvalid) O[i] = cv[i];
        } else {
            asm volatile("cp.async.wait_group 1;" ::: "memory");     // everything but the row just requested has landed
            __syncwarp();
            if (active) {
                const float* cb = capbuf + (band * 2 + ((y - yb) & 1)) * 512 + 8 * g;
                const float4 c0 = *(const float4*)cb, c1 = *(const float4*)(cb + 4);
                const uint4 r0 = *(const uint4*)(cb + 256), r1 = *(const uint4*)(cb + 260);
                const float crv[8] = { c0.x, c0.y, c0.z, c0.w, c1.x, c1.y, c1.z, c1.w };
                const uint32_t rwd[8] = { r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w };
                uint32_t wc[NW];
                cr_load_row<NW>(wc, tp + (size_t)(R + 1) * PITCH);   // centre row y
                const int16_t* drow = a.dr + (size_t)frame * a.dr_frame + (size_t)gy * a.dr_step + a.rrx;
                const int xr = gx0 - a.rrx;
                const int par0 = (int)(((uintptr_t)drow >> 1) & 1);  // which half of the fetched word holds element ro: (par0 + ro) & 1
                float dc[8];
                uint8_t m[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const int L = cr_elem<NW>(wc, RA + i);
                    const int ro = xr + i - (L >> 4);                                              // 331
                    const bool hit = (unsigned)ro < (unsigned)a.rw;                                // 332
                    const int Rv = ((par0 + ro) & 1) ? ((int)rwd[i] >> 16) : (int)(short)(rwd[i] & 0xffffu);
                    const bool cons = abs(L + Rv) < a.thresh;                                      // 334
                    const float cap = hit ? (cons ? crv[i] : 0.0f) : __int_as_float(0x7f800000);   // min(confL, confR) / 0 / keep (335-337)
                    m[i] = hit ? (cons ? 2 : 3) : 1;
                    cv[i] = __fmul_rn(255.0f, fminf(cv[i], cap));                                  // 209
                    dc[i] = __fmul_rn(cv[i], (float)L);                                            // 289-290
                }
                if (a.mask && frame == 0)
#pragma unroll
                    for (int i = 0; i < 8; ++i) if (i < nvalid) a.mask[(size_t)gy * a.mask_step + gx0 + i] = m[i];
            if (a.plane_c) {
                float* PC = (a.il ? a.plane_dc : a.plane_c) + plane_off(a.il, a.p_frame, a.pitch, frame, y, cx0, 1);
                float* PD = a.plane_dc + plane_off(a.il, a.p_frame, a.pitch, frame, y, cx0, 0);
                if (nvalid == 8) {
                    if (!a.skip_c) st_global_v8(PC, cv[0], cv[1], cv[2], cv[3], cv[4], cv[5], cv[6], cv[7]);
                    if (!a.skip_dc) st_global_v8(PD, dc[0], dc[1], dc[2], dc[3], dc[4], dc[5], dc[6], dc[7]);
                } else
#pragma unroll
                    for (int i = 0; i < 8; ++i) if (i < nvalid) { if (!a.skip_c) PC[i] = cv[i]; if (!a.skip_dc) PD[i] = dc[i]; }
            }
            if (a.conf) {
                float* E = a.conf + (size_t)frame * a.conf_frame + (size_t)gy * a.conf_step + gx0;
                if (nvalid == 8 && (((uintptr_t)E) & 31) == 0) {
                    st_global_v8(E, cv[0], cv[1], cv[2], cv[3], cv[4], cv[5], cv[6], cv[7]);
                } else
#pragma unroll
                    for (int i = 0; i < 8; ++i) if (i < nvalid) E[i] = cv[i];
            }
            }
            __syncwarp();                          // the buffer is rewritten by the next row's copies
        }
    }
}

}  // namespace dwls
