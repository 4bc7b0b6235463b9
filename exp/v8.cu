__global__ void k(float* o, const float* in)
{
    float v[8];
    for (int i = 0; i < 8; ++i) v[i] = in[threadIdx.x * 8 + i] * 2.f;
    float* p = o + threadIdx.x * 8;
    asm volatile("st.global.v8.f32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(p), "f"(v[0]), "f"(v[1]), "f"(v[2]), "f"(v[3]), "f"(v[4]), "f"(v[5]), "f"(v[6]), "f"(v[7]) : "memory");
}
