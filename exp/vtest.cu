#include <cuda_runtime.h>
__global__ void k(const unsigned* a, const unsigned* b, unsigned* o)
{
    unsigned x = a[threadIdx.x], y = b[threadIdx.x];
    unsigned d = __vabsdiffu4(x, y) & 0x00ffffffu;
    o[threadIdx.x] = __dp4a(d, d, 0u);
}
