#include <cuda.h>
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <stdint.h>
#include <string.h>
#include "../addingdisparityfiltering_b200/csrc/dwls_solver.cuh"
using namespace dwls;
__device__ __forceinline__ void tma3(void* dst, const void* tmap, int c0, int c1, int c2, uint64_t* bar)
{
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];" ::"r"(smem_u32(dst)),
                 "l"(tmap), "r"(c0), "r"(c1), "r"(c2), "r"(smem_u32(bar)) : "memory");
}
__global__ void k(const __grid_constant__ CUtensorMap tm, int bc, int br, int x, int y, int z, int* out)
{
    extern __shared__ __align__(128) uint16_t tile[];
    __shared__ __align__(8) uint64_t bar;
    if (threadIdx.x == 0) { mbar_init(&bar, 1); mbar_fence_init(); }
    __syncthreads();
    if (threadIdx.x == 0) { mbar_arrive_expect_tx(&bar, bc * br * 2); tma3(tile, &tm, x, y, z, &bar); }
    mbar_wait(&bar, 0);
    long s = 0;
    for (int i = threadIdx.x; i < bc * br; i += blockDim.x) s += (short)tile[i];
    atomicAdd(out, (int)s);
}
typedef CUresult (*Enc)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
int main(int argc, char** argv)
{
    int W = atoi(argv[1]), H = atoi(argv[2]), F = atoi(argv[3]), bc = atoi(argv[4]), br = atoi(argv[5]), x = atoi(argv[6]), y = atoi(argv[7]);
    int16_t* d; cudaMalloc(&d, (size_t)W * H * F * 2);
    int16_t* h = (int16_t*)malloc((size_t)W * H * F * 2);
    for (size_t i = 0; i < (size_t)W * H * F; ++i) h[i] = 1;
    cudaMemcpy(d, h, (size_t)W * H * F * 2, cudaMemcpyHostToDevice);
    void* fn; cudaDriverEntryPointQueryResult q;
    cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
    CUtensorMap tm;
    cuuint64_t dims[3] = { (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)F }, str[2] = { (cuuint64_t)W * 2, (cuuint64_t)W * H * 2 };
    cuuint32_t box[3] = { (cuuint32_t)bc, (cuuint32_t)br, 1 }, es[3] = { 1, 1, 1 };
    CUresult r = ((Enc)fn)(&tm, CU_TENSOR_MAP_DATA_TYPE_UINT16, 3, d, dims, str, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("encode=%d ", (int)r);
    int* out; cudaMalloc(&out, 4); cudaMemset(out, 0, 4);
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, bc * br * 2);
    k<<<1, 128, bc * br * 2>>>(tm, bc, br, x, y, 0, out);
    cudaError_t e = cudaDeviceSynchronize();
    int ho = 0; cudaMemcpy(&ho, out, 4, cudaMemcpyDeviceToHost);
    printf("W=%d H=%d F=%d box=%dx%d at (%d,%d): %s sum=%d\n", W, H, F, bc, br, x, y, cudaGetErrorString(e), ho);
    return 0;
}
