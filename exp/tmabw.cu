// TMA 2-D tile load throughput vs box width: each CTA streams boxes {bw floats x br rows} (bw*br*4 = 32 KB) from a big matrix
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <stdint.h>
#include "../addingdisparityfiltering_b200/csrc/dwls_solver.cuh"
using namespace dwls;
__global__ void __launch_bounds__(128) k(const __grid_constant__ CUtensorMap tm, int bw, int br, int boxes_x, int boxes_y, int nstage, float* out)
{
    extern __shared__ __align__(128) float sm[];
    __shared__ __align__(8) uint64_t bar[8];
    const int t = threadIdx.x;
    const int box_floats = bw * br;
    if (t == 0) { for (int i = 0; i < nstage; ++i) mbar_init(bar + i, 1); mbar_fence_init(); }
    __syncthreads();
    const int total = boxes_x * boxes_y;
    const int n_mine = (total - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
    auto issue = [&](int i) {
        const int b = blockIdx.x + i * gridDim.x, by = b / boxes_x, bx = b - by * boxes_x;
        mbar_arrive_expect_tx(bar + (i % nstage), box_floats * 4);
        for (int r = 0; r < br; r += 256)
            tma_load_2d(sm + (size_t)(i % nstage) * box_floats + (size_t)r * bw, &tm, bx * bw, by * br + r, bar + (i % nstage));
    };
    if (t == 0) for (int i = 0; i < nstage && i < n_mine; ++i) issue(i);
    float acc = 0.f;
    for (int i = 0; i < n_mine; ++i) {
        mbar_wait(bar + (i % nstage), (i / nstage) & 1);
        acc += sm[(size_t)(i % nstage) * box_floats + t];
        __syncthreads();
        if (t == 0 && i + nstage < n_mine) issue(i + nstage);
    }
    if (acc == 123.456f) out[0] = acc;
}
typedef CUresult (*Enc)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
int main(int argc, char** argv)
{
    const int W = 1760, H = 1080 * 32;                       // 32 frames of ROI planes
    float* d; cudaMalloc(&d, (size_t)W * H * 4); cudaMemset(d, 0, (size_t)W * H * 4);
    float* out; cudaMalloc(&out, 4);
    void* fn; cudaDriverEntryPointQueryResult q;
    cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
    const int bws[] = { 8, 16, 32, 80, 160 };
    for (int nstage = 1; nstage <= 2; ++nstage)
    for (int bi = 0; bi < 5; ++bi) {
        const int bw = bws[bi];
        int br = 1080; int boxr = 216;                       // tile = bw x 1080 rows fetched as 5 boxes of 216 rows
        if (bw * br * 4 * nstage > 200 * 1024) continue;
        CUtensorMap tm;
        cuuint64_t dims[2] = { (cuuint64_t)W, (cuuint64_t)H }, str[1] = { (cuuint64_t)W * 4 };
        cuuint32_t box[2] = { (cuuint32_t)bw, (cuuint32_t)boxr }, es[2] = { 1, 1 };
        CUresult r = ((Enc)fn)(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, d, dims, str, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r) { printf("encode fail %d\n", (int)r); continue; }
        // kernel issues boxes of 256 rows in its loop; use br = 1080 with boxr = 216 -> adapt: pass br=216*5 and step 216
        const size_t smem = (size_t)bw * br * 4 * nstage;
        cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
        for (int rep = 0; rep < 3; ++rep) {
            cudaEventRecord(e0);
            k<<<148, 128, smem>>>(tm, bw, 216, W / bw, H / 216, nstage * 5, out);   // stage = one 216-row box; 5*nstage boxes in flight
            cudaEventRecord(e1); cudaEventSynchronize(e1);
        }
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        printf("bw=%3d floats (%3d B rows) in-flight/SM=%6.1f KB : %.3f ms  %.0f GB/s  (%s)\n", bw, bw * 4, bw * 216 * 4 * nstage * 5 / 1024.0, ms,
               (double)W * H * 4 / ms / 1e6, cudaGetErrorString(cudaGetLastError()));
    }
    return 0;
}
